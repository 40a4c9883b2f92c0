// reduce.cuh -- deterministic grid-wide reductions (warp shuffle -> block -> last-block pass).
// Mirrors the reference's reduce=+ / min / max KERNELs (codegen_kernel.cpp:359-363), whose
// join order is thread-schedule dependent; here the order is fixed, so results are
// run-to-run bit-stable.
#pragma once
#include "common.cuh"

namespace mf {

constexpr int RED_THREADS = 256;
constexpr int RED_MAX_BLOCKS = 2048;

struct OpSum { __device__ static double id() { return 0.0; } __device__ static double f(double a, double b) { return a + b; } };
struct OpMin { __device__ static double id() { return 1e300; } __device__ static double f(double a, double b) { return a < b ? a : b; } };
struct OpMax { __device__ static double id() { return -1e300; } __device__ static double f(double a, double b) { return a > b ? a : b; } };

template <class Op>
__device__ __forceinline__ double blockReduce(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v = Op::f(v, __shfl_xor_sync(0xffffffffu, v, o));
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;   // 1-D blocks only
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? sh[threadIdx.x] : Op::id();
  if (w == 0) for (int o = 16; o > 0; o >>= 1) v = Op::f(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;  // valid in thread 0
}

// Finish a K-valued reduction: every block calls this with its K block-level values (valid in
// thread 0); the last block to arrive combines the per-block partials in block order and writes out[k].
template <class Op0, class Op1, int K>
__device__ __forceinline__ void gridFinish(double v0, double v1, double* partials, unsigned int* counter, double* out, double* sh) {
  __shared__ bool isLast;
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = v0;
    if (K > 1) partials[RED_MAX_BLOCKS + blockIdx.x] = v1;
    __threadfence();
    unsigned int t = atomicAdd(counter, 1u);
    isLast = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (isLast) {
    __threadfence();
    double a = Op0::id(), b = Op1::id();
    for (unsigned int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
      a = Op0::f(a, ((volatile double*)partials)[i]);
      if (K > 1) b = Op1::f(b, ((volatile double*)partials)[RED_MAX_BLOCKS + i]);
    }
    a = blockReduce<Op0>(a, sh);
    if (K > 1) b = blockReduce<Op1>(b, sh);
    if (threadIdx.x == 0) { out[0] = a; if (K > 1) out[1] = b; *counter = 0; }
  }
}

// ---- K-valued variant with run-time ops (0 = sum, 1 = min, 2 = max) and a finishing functor that runs in
// thread 0 of the last block with the K grid-wide results (used by the CG kernels to update device-side scalars).
__device__ __forceinline__ double redIdent(int op) { return op == 0 ? 0.0 : (op == 1 ? 1e300 : -1e300); }
__device__ __forceinline__ double redApply(int op, double a, double b) { return op == 0 ? a + b : (op == 1 ? (a < b ? a : b) : (a > b ? a : b)); }
__device__ __forceinline__ double blockReduceOp(double v, int op, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v = redApply(op, v, __shfl_xor_sync(0xffffffffu, v, o));
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? sh[threadIdx.x] : redIdent(op);
  if (w == 0) for (int o = 16; o > 0; o >>= 1) v = redApply(op, v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
template <int K, class Fin>
__device__ __forceinline__ void gridReduceK(double (&v)[K], const int (&ops)[K], double* partials, unsigned int* counter, double* sh, Fin fin) {
  __shared__ bool isLastK;
#pragma unroll
  for (int q = 0; q < K; q++) v[q] = blockReduceOp(v[q], ops[q], sh);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int q = 0; q < K; q++) partials[q * RED_MAX_BLOCKS + blockIdx.x] = v[q];
    __threadfence();
    const unsigned int t = atomicAdd(counter, 1u);
    isLastK = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (isLastK) {
    __threadfence();
    double r[K];
#pragma unroll
    for (int q = 0; q < K; q++) {
      double a = redIdent(ops[q]);
      for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x) a = redApply(ops[q], a, ((volatile double*)partials)[q * RED_MAX_BLOCKS + b]);
      r[q] = blockReduceOp(a, ops[q], sh);
    }
    if (threadIdx.x == 0) { *counter = 0; fin(r); }
  }
}

inline int redBlocks(const Solver* s, long long n) {
  long long b = (n + RED_THREADS * 4 - 1) / (RED_THREADS * 4);
  long long cap = (long long)s->numSMs * 8;
  if (cap > RED_MAX_BLOCKS) cap = RED_MAX_BLOCKS;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// fetch `count` doubles of s->redOut to the host (synchronises the solver stream)
void fetchRed(Solver* s, int count);

// ---- persistent tile loop for ijk kernels that also reduce ---------------------------------------
// 256-thread 1-D blocks walk 32x4x2 (3-D) or 32x8x1 (2-D) cell tiles in a grid-stride loop, so the
// number of per-block partials stays bounded (<= RED_MAX_BLOCKS) at any grid size.
#define MF_TILE_LOOP_BEGIN(g) \
  { const int _ty = (g).is3D ? 4 : 8, _tz = (g).is3D ? 2 : 1; \
    const int _ntx = ((g).nx + 31) >> 5, _nty = ((g).ny + _ty - 1) / _ty, _ntz = ((g).nz + _tz - 1) / _tz; \
    const long long _nt = (long long)_ntx * _nty * _ntz; \
    const int _lx = threadIdx.x & 31; \
    const int _ly = (g).is3D ? ((threadIdx.x >> 5) & 3) : (threadIdx.x >> 5); \
    const int _lz = (g).is3D ? (threadIdx.x >> 7) : 0; \
    for (long long _t = blockIdx.x; _t < _nt; _t += gridDim.x) { \
      const int _bx = (int)(_t % _ntx), _by = (int)((_t / _ntx) % _nty), _bz = (int)(_t / ((long long)_ntx * _nty)); \
      const int i = _bx * 32 + _lx, j = _by * _ty + _ly, k = _bz * _tz + _lz; \
      if (i < (g).nx && j < (g).ny && k < (g).nz) { \
        const long long idx = (long long)i + (long long)(g).nx * j + (long long)(g).nx * (g).ny * k;
#define MF_TILE_LOOP_END }}}

inline int tileBlocks(const Solver* s) {
  const int ty = s->is3D ? 4 : 8, tz = s->is3D ? 2 : 1;
  long long nt = (long long)((s->nx + 31) / 32) * ((s->ny + ty - 1) / ty) * ((s->nz + tz - 1) / tz);
  long long cap = (long long)s->numSMs * 8;
  if (cap > RED_MAX_BLOCKS) cap = RED_MAX_BLOCKS;
  return (int)(nt < cap ? nt : cap);
}

} // namespace mf
