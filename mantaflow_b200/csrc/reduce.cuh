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

// ---- K-valued reductions with run-time ops (0 = sum, 1 = min, 2 = max) and a finishing functor ---------------------
// Single GPU: the last block of the reducing kernel combines the per-block partials in block order and runs fin(r) in its
// thread 0 (fin updates device-resident scalars, e.g. the CG state).
// Several ranks (z-slabs on several GPUs, or emulated on one): the last block instead PUSHES the rank's K results into slot
// red[parity][rank][] of EVERY rank's control block (peer stores over NVLink); after the cross-rank barrier each process runs
// k_red_finish, which combines the slots in rank order and runs fin -- every rank computes bit-identical scalars.
struct RedCtx {
  double* partials; unsigned int* counter;
  int nranks, rank;
  char* shared; size_t stride;     // CommShared blocks (comm.h)
  int parity, self;
};
template <int K> struct RedOps { int op[K]; };

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
__device__ __forceinline__ double* redSlot(const RedCtx& c, int host, int rank) {
  // CommShared::red[parity][rank][k], after CommShared::flag[COMM_MAX_RANKS] (comm.h)
  return reinterpret_cast<double*>(c.shared + (size_t)host * c.stride + 16 * sizeof(unsigned long long)) + ((size_t)c.parity * 16 + rank) * 8;
}
template <int K, class Fin>
__device__ __forceinline__ void gridReduceK(double (&v)[K], const int (&ops)[K], const RedCtx& ctx, double* sh, Fin fin) {
  __shared__ bool isLastK;
  double* partials = ctx.partials;
#pragma unroll
  for (int q = 0; q < K; q++) v[q] = blockReduceOp(v[q], ops[q], sh);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int q = 0; q < K; q++) partials[q * RED_MAX_BLOCKS + blockIdx.x] = v[q];
    __threadfence();
    const unsigned int t = atomicAdd(ctx.counter, 1u);
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
    if (threadIdx.x == 0) {
      *ctx.counter = 0;
      if (ctx.nranks == 1) fin(r);
      else {
        for (int h = 0; h < ctx.nranks; h++) {
          double* slot = redSlot(ctx, h, ctx.rank);
#pragma unroll
          for (int q = 0; q < K; q++) slot[q] = r[q];
        }
      }
    }
  }
}
template <int K, class Fin>
__global__ void k_red_finish(RedOps<K> ops, RedCtx ctx, Fin fin) {
  MF_PDL_PROLOGUE
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double r[K];
#pragma unroll
  for (int q = 0; q < K; q++) {
    double a = redIdent(ops.op[q]);
    for (int h = 0; h < ctx.nranks; h++) a = redApply(ops.op[q], a, ((volatile double*)redSlot(ctx, ctx.self, h))[q]);
    r[q] = a;
  }
  fin(r);
}
// r[k] -> out[k]
template <int K> struct FinStore { double* out; __device__ void operator()(double* r) const { for (int q = 0; q < K; q++) out[q] = r[q]; } };

RedCtx redCtx(Solver* s, int rank);
// after the per-rank launches of a reducing kernel: barrier, then combine the ranks' slots and run fin on this process
template <int K, class Fin>
inline void redFinish(Solver* s, const RedOps<K>& ops, Fin fin) {
  if (!s->multi()) return;
  s->rankSync();
  RedCtx c = redCtx(s, s->rankBegin);
  launchPdl(s->stream, k_red_finish<K, Fin>, dim3(1), dim3(32), ops, c, fin);
  s->launches++;
  s->redParity ^= 1;
}

// perSM: CTAs of the kernel one SM holds (grid-stride kernels: a grid of exactly the resident CTAs has no partial last wave)
inline int redBlocks(const Solver* s, long long n, int perSM = 8) {
  long long b = (n + RED_THREADS * 4 - 1) / (RED_THREADS * 4);
  long long cap = (long long)s->numSMs * perSM;
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
    const int _ntx = ((g).nx + 31) >> 5, _nty = ((g).ny + _ty - 1) / _ty, _ntz = ((g).k1 - (g).k0 + _tz - 1) / _tz; \
    const long long _nt = (long long)_ntx * _nty * _ntz; \
    const int _lx = threadIdx.x & 31; \
    const int _ly = (g).is3D ? ((threadIdx.x >> 5) & 3) : (threadIdx.x >> 5); \
    const int _lz = (g).is3D ? (threadIdx.x >> 7) : 0; \
    for (long long _t = blockIdx.x; _t < _nt; _t += gridDim.x) { \
      const int _bx = (int)(_t % _ntx), _by = (int)((_t / _ntx) % _nty), _bz = (int)(_t / ((long long)_ntx * _nty)); \
      const int i = _bx * 32 + _lx, j = _by * _ty + _ly, k = (g).k0 + _bz * _tz + _lz; \
      if (i < (g).nx && j < (g).ny && k < (g).k1) { \
        const long long idx = (long long)i + (long long)(g).nx * j + (long long)(g).nx * (g).ny * k;
#define MF_TILE_LOOP_END }}}

// The same with W x-consecutive cells per thread (quad.cuh): tiles of 32 threads x 4 x 2 rows; a warp is one row, so the row
// test is warp-uniform and out-of-range lanes stay (clamped to the last cells of the row, on == false) for the shuffles.
#define MF_QTILE_LOOP_BEGIN(g, W) \
  { const int _ty = (g).is3D ? 4 : 8, _tz = (g).is3D ? 2 : 1; \
    const unsigned _ntx = (unsigned)((((g).nx + (W) - 1) / (W) + 31) >> 5), _nty = (unsigned)(((g).ny + _ty - 1) / _ty); \
    const unsigned _nt = _ntx * _nty * (unsigned)(((g).k1 - (g).k0 + _tz - 1) / _tz); \
    const int _lx = threadIdx.x & 31; \
    const int _ly = (g).is3D ? ((threadIdx.x >> 5) & 3) : (threadIdx.x >> 5); \
    const int _lz = (g).is3D ? (threadIdx.x >> 7) : 0; \
    for (unsigned _t = blockIdx.x; _t < _nt; _t += gridDim.x) { \
      const unsigned _bx = _t % _ntx, _rr = _t / _ntx, _by = _rr % _nty, _bz = _rr / _nty; \
      const int _i0 = (int)(_bx * 32 + _lx) * (W), j = (int)_by * _ty + _ly, k = (g).k0 + (int)_bz * _tz + _lz; \
      if (j < (g).ny && k < (g).k1) { \
        const bool on = _i0 < (g).nx; \
        const int i = on ? _i0 : (g).nx - (W); \
        const long long idx = (long long)i + (long long)(g).nx * j + (long long)(g).nx * (g).ny * k;
#define MF_QTILE_LOOP_END }}}

inline int quadTileBlocks(const Solver* s, const Geo& g, int W) {
  const int ty = s->is3D ? 4 : 8, tz = s->is3D ? 2 : 1;
  long long nt = (long long)(((s->nx + W - 1) / W + 31) / 32) * ((s->ny + ty - 1) / ty) * ((g.k1 - g.k0 + tz - 1) / tz);
  if (nt < 1) nt = 1;
  long long cap = (long long)s->numSMs * 8;
  if (cap > RED_MAX_BLOCKS) cap = RED_MAX_BLOCKS;
  return (int)(nt < cap ? nt : cap);
}

inline int tileBlocks(const Solver* s, const Geo& g) {
  const int ty = s->is3D ? 4 : 8, tz = s->is3D ? 2 : 1;
  long long nt = (long long)((s->nx + 31) / 32) * ((s->ny + ty - 1) / ty) * ((g.k1 - g.k0 + tz - 1) / tz);
  if (nt < 1) nt = 1;
  long long cap = (long long)s->numSMs * 8;
  if (cap > RED_MAX_BLOCKS) cap = RED_MAX_BLOCKS;
  return (int)(nt < cap ? nt : cap);
}

} // namespace mf
