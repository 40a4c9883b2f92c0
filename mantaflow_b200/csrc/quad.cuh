// quad.cuh -- W x-consecutive cells per thread for the streaming ijk kernels (W = 4 when nx % 4 == 0, else 1).
// With W = 4 every array access is one 16-byte (8-byte for 16-bit data) transaction per thread and the index arithmetic is
// paid once per four cells; the -x / +x neighbours come from the adjacent lane by shuffle.  A warp is always 32 consecutive
// threads of ONE grid row (blockDim.x == 32), so out-of-range rows leave warp-uniformly and out-of-range lanes of a row stay for
// the shuffles (they compute on a clamped, valid position and store nothing).
#pragma once
#include "common.cuh"

namespace mf {

// i, j, k, idx of the thread's first cell; `on` = the cells exist.  Leaves (warp-uniformly) when the row does not exist.
#define MF_CELLS_W(g, W) \
  const int j = blockIdx.y * blockDim.y + threadIdx.y; \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + (g).k0; \
  if (j >= (g).ny || k >= (g).k1) return; \
  const int _i0 = (blockIdx.x * blockDim.x + threadIdx.x) * (W); \
  const bool on = _i0 < (g).nx; \
  const int i = on ? _i0 : (g).nx - (W); \
  const long long idx = (long long)i + (long long)(g).nx * j + (long long)(g).nx * (g).ny * k;

template <int W> __device__ __forceinline__ void ldW(float (&o)[W], const float* __restrict__ p, long long idx) {
  if constexpr (W == 4) { const float4 v = *reinterpret_cast<const float4*>(p + idx); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
  else { for (int c = 0; c < W; c++) o[c] = p[idx + c]; }
}
template <int W> __device__ __forceinline__ void ldW(int (&o)[W], const int* __restrict__ p, long long idx) {
  if constexpr (W == 4) { const int4 v = *reinterpret_cast<const int4*>(p + idx); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
  else { for (int c = 0; c < W; c++) o[c] = p[idx + c]; }
}
template <int W> __device__ __forceinline__ void stW(float* __restrict__ p, long long idx, const float (&v)[W]) {
  if constexpr (W == 4) *reinterpret_cast<float4*>(p + idx) = make_float4(v[0], v[1], v[2], v[3]);
  else { for (int c = 0; c < W; c++) p[idx + c] = v[c]; }
}
template <int W> __device__ __forceinline__ void stW(unsigned short* __restrict__ p, long long idx, const int (&v)[W]) {
  if constexpr (W == 4) {
    uint2 u; u.x = (unsigned)(v[0] & 0xffff) | ((unsigned)v[1] << 16); u.y = (unsigned)(v[2] & 0xffff) | ((unsigned)v[3] << 16);
    *reinterpret_cast<uint2*>(p + idx) = u;
  } else { for (int c = 0; c < W; c++) p[idx + c] = (unsigned short)v[c]; }
}

// o[c] = value of the cell left of cell c (x - 1); `edge` where that cell does not exist (i == 0)
template <int W, class T>
__device__ __forceinline__ void xMinusW(T (&o)[W], const T (&c)[W], const T* __restrict__ p, long long idx, int i, T edge) {
  if constexpr (W == 1) { o[0] = (i > 0) ? p[idx - 1] : edge; return; }
  T left = __shfl_up_sync(0xffffffffu, c[W - 1], 1);
  if ((threadIdx.x & 31) == 0) left = (i > 0) ? p[idx - 1] : edge;
  o[0] = left;
#pragma unroll
  for (int q = 1; q < W; q++) o[q] = c[q - 1];
}
// o[c] = value of the cell right of cell c (x + 1); `edge` beyond the row
template <int W, class T>
__device__ __forceinline__ void xPlusW(T (&o)[W], const T (&c)[W], const T* __restrict__ p, long long idx, int i, int nx, T edge) {
  if constexpr (W == 1) { o[0] = (i + 1 < nx) ? p[idx + 1] : edge; return; }
  T right = __shfl_down_sync(0xffffffffu, c[0], 1);
  if ((threadIdx.x & 31) == 31 || i + 2 * W > nx) right = (i + W < nx) ? p[idx + W] : edge;
#pragma unroll
  for (int q = 0; q + 1 < W; q++) o[q] = c[q + 1];
  o[W - 1] = right;
}

// The same in two steps, for kernels that issue ALL their loads before the first use: warps issue in order, so a shuffle placed
// between the loads would wait for its operand and hold the loads behind it back (one round trip to memory per group of loads
// instead of one per thread -- k_add_buoyancy / k_correct_velocity / k_set_wall_bcs were latency-bound that way).
// xEdgeMinus / xEdgePlus: the one value per warp that no lane holds (lane 0's left cell, lane 31's right cell), loaded with
// the other loads; xMinusWS / xPlusWS: the shuffle, after the loads.
template <int W, class T>
__device__ __forceinline__ T xEdgeMinus(const T* __restrict__ p, long long idx, int i, T edge) {
  return ((threadIdx.x & 31) == 0 && i > 0) ? p[idx - 1] : edge;
}
template <int W, class T>
__device__ __forceinline__ T xEdgePlus(const T* __restrict__ p, long long idx, int i, int nx, T edge) {
  return (((threadIdx.x & 31) == 31 || i + 2 * W > nx) && i + W < nx) ? p[idx + W] : edge;
}
template <int W, class T>
__device__ __forceinline__ void xMinusWS(T (&o)[W], const T (&c)[W], T le) {
  static_assert(W > 1, "one cell per thread: use xMinusW");
  T left = __shfl_up_sync(0xffffffffu, c[W - 1], 1);
  if ((threadIdx.x & 31) == 0) left = le;
  o[0] = left;
#pragma unroll
  for (int q = 1; q < W; q++) o[q] = c[q - 1];
}
template <int W, class T>
__device__ __forceinline__ void xPlusWS(T (&o)[W], const T (&c)[W], T re, int i, int nx) {
  static_assert(W > 1, "one cell per thread: use xPlusW");
  T right = __shfl_down_sync(0xffffffffu, c[0], 1);
  if ((threadIdx.x & 31) == 31 || i + 2 * W > nx) right = re;
#pragma unroll
  for (int q = 0; q + 1 < W; q++) o[q] = c[q + 1];
  o[W - 1] = right;
}

inline dim3 quadGridOf(dim3 b, const Geo& g, int W) { return dim3(cdiv(cdiv(g.nx, W), b.x), cdiv(g.ny, b.y), cdiv(g.k1 - g.k0, b.z)); }
// ijk kernel template kern<W> over every local rank's z range (first parameter Geo), then the cross-rank barrier
#define MF_LAUNCH_IJK_W(s, kern, ...) do { const dim3 _b = mf::ijkBlockOf(s); MF_RANKS(s, _r) { const mf::Geo _g = mf::geo((s), _r); \
  if (_g.k1 > _g.k0) { if ((s)->nx % 4 == 0) { MF_LAUNCH(s, kern<4>, mf::quadGridOf(_b, _g, 4), _b, _g, __VA_ARGS__); } \
                       else { MF_LAUNCH(s, kern<1>, mf::quadGridOf(_b, _g, 1), _b, _g, __VA_ARGS__); } } } (s)->rankSync(); } while (0)

// the same with programmatic dependent launch: for kernels that begin with MF_PDL_PROLOGUE
#define MF_LAUNCH_IJK_W_PDL(s, kern, ...) do { const dim3 _b = mf::ijkBlockOf(s); MF_RANKS(s, _r) { const mf::Geo _g = mf::geo((s), _r); \
  if (_g.k1 > _g.k0) { if ((s)->nx % 4 == 0) { MF_LAUNCH_STEP(s, kern<4>, mf::quadGridOf(_b, _g, 4), _b, _g, __VA_ARGS__); } \
                       else { MF_LAUNCH_STEP(s, kern<1>, mf::quadGridOf(_b, _g, 1), _b, _g, __VA_ARGS__); } } } (s)->rankSync(); } while (0)

}  // namespace mf
