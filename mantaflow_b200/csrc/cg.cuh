// cg.cuh -- preconditioned conjugate gradient on the 7-point stencil, scalars resident on the device.
// Reference: GridCg<APPLYMAT> (source/conjugategrad.h:65-114, conjugategrad.cpp:201-335):
//   doInit :210, iterate :238 (ApplyMatrix -> GridDotProduct -> 2x gridScaledAdd -> preconditioner ->
//   getMaxAbs | GridSumSqr -> GridDotProduct -> UpdateSearchVec).
// The reference makes >= 9 full-grid sweeps and 4 host-visible reductions per iteration. Here one
// iteration is 3 kernels (+ the preconditioner): alpha, beta, the residual norm and the convergence flag
// are produced by the last block of the reducing kernel and consumed from device memory by the next
// kernel, so the host only polls the flag.  Arithmetic (float products, double accumulation, float
// scalars) follows the reference; only the summation order of the reductions differs (fixed here).
#pragma once
#include "common.cuh"
#include "reduce.cuh"
#include "pressure_kernels.cuh"

namespace mf {

struct CgState {        // lives in device memory
  float sigma;          // mSigma
  float alpha, beta;
  float resNorm;        // mResNorm
  float accuracy;       // mAccuracy
  int useL2;
  int done;             // iterate() returned false
  int diverged;
  int iterations;       // mIterations
  int pad;
};

struct CgVecs {
  const int* flags;
  float* dst; float* residual; float* search; float* tmp;
  const float* rhs;
  const float *A0, *Ai, *Aj, *Ak;
  const unsigned short* code;   // 16-bit row codes (pressure_kernels.cuh CGC_*), or null: use the float planes
};

// finishing functors of the CG reductions: run once per process with the grid-wide (and rank-wide) results
struct FinAlpha {   // dp = tmp . search ; alpha = sigma / dp (conjugategrad.cpp:250-252)
  CgState* st;
  __device__ void operator()(double* r) const {
    if (st->done) return;
    const float dp = (float)r[0];
    float alpha = 0.f;
    if (fabsf(dp) > 0.f) alpha = st->sigma / dp;
    st->alpha = alpha;
    st->iterations += 1;
  }
};
template <bool PCNONE> struct FinUpdate {   // residual norm and stop test (conjugategrad.cpp:262-277) [+ sigma, beta when unpreconditioned]
  CgState* st;
  __device__ void operator()(double* r) const {
    if (st->done) return;
    float resNorm;
    if (st->useL2) resNorm = (float)r[0];
    else resNorm = fmaxf(fabsf((float)r[1]), fabsf((float)r[2]));
    st->resNorm = resNorm;
    if (resNorm < st->accuracy) { st->sigma = resNorm; st->done = 1; return; }
    if (PCNONE) {
      const float sigmaNew = (float)r[3];
      st->beta = sigmaNew / st->sigma;
      st->sigma = sigmaNew;
      if (!(resNorm < 1e35f)) { st->diverged = 1; st->done = 1; }
    }
  }
};
struct FinBeta {    // sigmaNew, beta = sigmaNew / sigma (conjugategrad.cpp:279-295)
  CgState* st;
  __device__ void operator()(double* r) const {
    if (st->done) return;
    const float sigmaNew = (float)r[0];
    st->beta = sigmaNew / st->sigma;
    st->sigma = sigmaNew;
    if (!(st->resNorm < 1e35f)) { st->diverged = 1; st->done = 1; }
  }
};
struct FinSigma { CgState* st; __device__ void operator()(double* r) const { st->sigma = (float)r[0]; } };

// tmp = A * search ; dp = tmp . search ; alpha = sigma / dp        (conjugategrad.cpp:247-252)
__global__ void __launch_bounds__(256) k_cg_spmv_dot(Geo g, CgVecs v, CgState* st, RedCtx ctx) {
  MF_PDL_PROLOGUE
  __shared__ double sh[32];
  if (st->done) return;
  double acc[1] = { 0.0 };
  MF_TILE_LOOP_BEGIN(g)
    const float s = v.search[idx];
    const float t = fIsFluid(v.flags[idx]) ? applyRow(g, idx, v.search, v.A0, v.Ai, v.Aj, v.Ak) : s;
    v.tmp[idx] = t;
    acc[0] += (double)(t * s);
  MF_TILE_LOOP_END
  const int ops[1] = { 0 };
  gridReduceK<1>(acc, ops, ctx, sh, FinAlpha{ st });
}

// The same with the 16-bit row codes, four x-consecutive cells per thread (needs nx % 4 == 0).  ApplyMatrix is a pure
// linear-index stencil (conjugategrad.h:205-222), so the kernel is 1-D over quads of the rank's slab: one 8-byte code load, five
// 16-byte loads of search (centre, +-Y, +-Z), the two x neighbours by shuffle, one 16-byte store: 10 B/cell of DRAM traffic.
// The products use the reconstructed coefficients (-1 or 0) in the reference's order, so tmp is bit-identical to applyRow's.
__device__ __forceinline__ float rowFromCode(int c, float s, float xm, float xp, float ym, float yp, float zm, float zp, bool is3D) {
  float r = s * (float)((c >> CGC_DIAG_SHIFT) & 7);
  r += xm * ((c & CGC_XM) ? -1.0f : 0.0f);
  r += xp * ((c & CGC_XP) ? -1.0f : 0.0f);
  r += ym * ((c & CGC_YM) ? -1.0f : 0.0f);
  r += yp * ((c & CGC_YP) ? -1.0f : 0.0f);
  if (is3D) {
    r += zm * ((c & CGC_ZM) ? -1.0f : 0.0f);
    r += zp * ((c & CGC_ZP) ? -1.0f : 0.0f);
  }
  return (c & CGC_FLUID) ? r : s;
}
__global__ void __launch_bounds__(256) k_cg_spmv_dot_code(Geo g, Span sp, CgVecs v, CgState* st, RedCtx ctx) {
  MF_PDL_PROLOGUE
  __shared__ double sh[32];
  if (st->done) return;
  double acc[1] = { 0.0 };
  const long long Y = g.nx, Z = g.Z, N = g.N;
  const long long q0 = sp.off >> 2, q1 = (sp.off + sp.n) >> 2;
  const int lane = threadIdx.x & 31;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  // whole warps walk the range so that the shuffles below are always executed by all 32 lanes
  for (long long qb = q0 + (blockIdx.x * (long long)blockDim.x + (threadIdx.x & ~31)); qb < q1; qb += (long long)gridDim.x * blockDim.x) {
    const long long q = qb + lane;
    const bool on = q < q1;
    const long long idx = (on ? q : q1 - 1) << 2;
    const float4 s = *reinterpret_cast<const float4*>(v.search + idx);
    const uint2 cw = *reinterpret_cast<const uint2*>(v.code + idx);
    const float4 ym = (idx - Y >= 0) ? *reinterpret_cast<const float4*>(v.search + idx - Y) : zero4;
    const float4 yp = (idx + Y < N) ? *reinterpret_cast<const float4*>(v.search + idx + Y) : zero4;
    float4 zm = zero4, zp = zero4;
    if (g.is3D) {
      if (idx - Z >= 0) zm = *reinterpret_cast<const float4*>(v.search + idx - Z);
      if (idx + Z < N) zp = *reinterpret_cast<const float4*>(v.search + idx + Z);
    }
    // the two x neighbours no lane of the warp holds are loaded with the other loads, before the shuffles wait for `s`
    const bool edgeM = lane == 0, edgeP = lane == 31 || q + 1 >= q1;
    const float xmE = (edgeM && idx - 1 >= 0) ? v.search[idx - 1] : 0.f;
    const float xpE = (edgeP && idx + 4 < N) ? v.search[idx + 4] : 0.f;
    float xm = __shfl_up_sync(0xffffffffu, s.w, 1), xp = __shfl_down_sync(0xffffffffu, s.x, 1);
    if (edgeM) xm = xmE;
    if (edgeP) xp = xpE;
    float4 t;
    t.x = rowFromCode((int)(cw.x & 0xffffu), s.x, xm, s.y, ym.x, yp.x, zm.x, zp.x, g.is3D);
    t.y = rowFromCode((int)(cw.x >> 16), s.y, s.x, s.z, ym.y, yp.y, zm.y, zp.y, g.is3D);
    t.z = rowFromCode((int)(cw.y & 0xffffu), s.z, s.y, s.w, ym.z, yp.z, zm.z, zp.z, g.is3D);
    t.w = rowFromCode((int)(cw.y >> 16), s.w, s.z, xp, ym.w, yp.w, zm.w, zp.w, g.is3D);
    if (on) {
      *reinterpret_cast<float4*>(v.tmp + idx) = t;
      acc[0] += (double)(t.x * s.x); acc[0] += (double)(t.y * s.y); acc[0] += (double)(t.z * s.z); acc[0] += (double)(t.w * s.w);
    }
  }
  const int ops[1] = { 0 };
  gridReduceK<1>(acc, ops, ctx, sh, FinAlpha{ st });
}

// The same product as a z-marching register pipeline (3-D): a warp owns 32 x-consecutive quads of one row and
// walks a chunk of z planes; the quads of planes k - 1, k, k + 1 of its column stay in registers, so every value of `search` is
// loaded once as a centre (plus the two row neighbours, which the neighbouring warps of the CTA have just pulled into L1) instead of
// five times.  Identical products in identical order per cell: tmp is bit-identical to k_cg_spmv_dot_code's.
// Work item = (row j, 32-quad segment of the row); a CTA takes 8 consecutive items, blockIdx.y the z chunk.
__global__ void __launch_bounds__(256) k_cg_spmv_dot_zm(Geo g, CgVecs v, CgState* st, RedCtx ctx, int zc, int ntiles, int units) {
  MF_PDL_PROLOGUE
  __shared__ double sh[32];
  if (st->done) return;
  double acc[1] = { 0.0 };
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nqx = g.nx >> 2, segs = (nqx + 31) >> 5;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int u = blockIdx.x; u < units; u += gridDim.x) {       // unit = (8 work items, z chunk); consecutive CTAs take consecutive tiles of a chunk
  const int item = (u % ntiles) * 8 + warp;
  const int j = item / segs, q = (item - j * segs) * 32 + lane;
  const int ka = g.k0 + (u / ntiles) * zc, kb = min(ka + zc, g.k1);
  if (j < g.ny && ka < kb) {                      // warp-uniform
    const bool on = q < nqx;
    const long long Y = g.nx, Z = g.Z, N = g.N;
    const long long col = (long long)4 * (on ? q : nqx - 1) + Y * j;
    const bool edgeM = lane == 0, edgeP = lane == 31 || q + 1 >= nqx;
    long long idx = col + Z * ka;
    float4 sm = (idx - Z >= 0) ? *reinterpret_cast<const float4*>(v.search + idx - Z) : zero4;
    float4 s0 = *reinterpret_cast<const float4*>(v.search + idx);
    for (int k = ka; k < kb; k++, idx += Z) {
      const float4 sp = (idx + Z < N) ? *reinterpret_cast<const float4*>(v.search + idx + Z) : zero4;
      const uint2 cw = *reinterpret_cast<const uint2*>(v.code + idx);
      const float4 ym = (idx - Y >= 0) ? *reinterpret_cast<const float4*>(v.search + idx - Y) : zero4;
      const float4 yp = (idx + Y < N) ? *reinterpret_cast<const float4*>(v.search + idx + Y) : zero4;
      const float xmE = (edgeM && idx - 1 >= 0) ? v.search[idx - 1] : 0.f;
      const float xpE = (edgeP && idx + 4 < N) ? v.search[idx + 4] : 0.f;
      float xm = __shfl_up_sync(0xffffffffu, s0.w, 1), xp = __shfl_down_sync(0xffffffffu, s0.x, 1);
      if (edgeM) xm = xmE;
      if (edgeP) xp = xpE;
      float4 t;
      t.x = rowFromCode((int)(cw.x & 0xffffu), s0.x, xm, s0.y, ym.x, yp.x, sm.x, sp.x, true);
      t.y = rowFromCode((int)(cw.x >> 16), s0.y, s0.x, s0.z, ym.y, yp.y, sm.y, sp.y, true);
      t.z = rowFromCode((int)(cw.y & 0xffffu), s0.z, s0.y, s0.w, ym.z, yp.z, sm.z, sp.z, true);
      t.w = rowFromCode((int)(cw.y >> 16), s0.w, s0.z, xp, ym.w, yp.w, sm.w, sp.w, true);
      if (on) {
        *reinterpret_cast<float4*>(v.tmp + idx) = t;
        acc[0] += (double)(t.x * s0.x); acc[0] += (double)(t.y * s0.y); acc[0] += (double)(t.z * s0.z); acc[0] += (double)(t.w * s0.w);
      }
      sm = s0; s0 = sp;
    }
  }
  }
  const int ops[1] = { 0 };
  gridReduceK<1>(acc, ops, ctx, sh, FinAlpha{ st });
}

// dst += alpha*search ; residual += -alpha*tmp ; norm(residual)     (conjugategrad.cpp:254-277)
// PCNONE: also tmp := residual (implicitly) and sigmaNew = residual . residual, beta        (:279-286)
template <bool PCNONE>
__device__ __forceinline__ float cgUpdateCell(float& d, float se, float r0, float t, float alpha, float nalpha, int useL2, double* acc) {
  d += alpha * se;
  const float r = r0 + nalpha * t;
  const double rd = (double)r;
  if (useL2) acc[0] += rd * rd;
  else { if (rd < acc[1]) acc[1] = rd; if (rd > acc[2]) acc[2] = rd; }
  if (PCNONE) acc[3] += (double)(r * r);
  return r;
}
// V4: the span starts and ends on a multiple of four cells (always the case when nx % 4 == 0): 16-byte accesses
template <bool PCNONE, bool V4>
__global__ void __launch_bounds__(256) k_cg_update(Span sp, CgVecs v, CgState* st, RedCtx ctx) {
  MF_PDL_PROLOGUE
  __shared__ double sh[32];
  if (st->done) return;
  const float alpha = st->alpha, nalpha = -alpha;
  const int useL2 = st->useL2;
  double acc[4] = { 0.0, 1e300, -1e300, 0.0 };   // sumsqr, min, max, r.r (float products)
  const long long st_ = (long long)gridDim.x * blockDim.x;
  if (V4) {
    const long long q1 = (sp.off + sp.n) >> 2;
    for (long long q = (sp.off >> 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; q < q1; q += st_) {
      float4 d = reinterpret_cast<float4*>(v.dst)[q];
      const float4 se = reinterpret_cast<const float4*>(v.search)[q];
      float4 r = reinterpret_cast<float4*>(v.residual)[q];
      const float4 t = reinterpret_cast<const float4*>(v.tmp)[q];
      r.x = cgUpdateCell<PCNONE>(d.x, se.x, r.x, t.x, alpha, nalpha, useL2, acc);
      r.y = cgUpdateCell<PCNONE>(d.y, se.y, r.y, t.y, alpha, nalpha, useL2, acc);
      r.z = cgUpdateCell<PCNONE>(d.z, se.z, r.z, t.z, alpha, nalpha, useL2, acc);
      r.w = cgUpdateCell<PCNONE>(d.w, se.w, r.w, t.w, alpha, nalpha, useL2, acc);
      reinterpret_cast<float4*>(v.dst)[q] = d;
      reinterpret_cast<float4*>(v.residual)[q] = r;
    }
  } else {
    const long long i0 = sp.off + blockIdx.x * (long long)blockDim.x + threadIdx.x, iend = sp.off + sp.n;
    for (long long idx = i0; idx < iend; idx += st_) {
      float d = v.dst[idx];
      v.residual[idx] = cgUpdateCell<PCNONE>(d, v.search[idx], v.residual[idx], v.tmp[idx], alpha, nalpha, useL2, acc);
      v.dst[idx] = d;
    }
  }
  const int ops[4] = { 0, 1, 2, 0 };
  gridReduceK<4>(acc, ops, ctx, sh, FinUpdate<PCNONE>{ st });
}

// sigmaNew = tmp . residual ; beta = sigmaNew / sigma                 (conjugategrad.cpp:279-295), preconditioned path
template <bool V4>
__global__ void __launch_bounds__(256) k_cg_dot_beta(long long n, const float* __restrict__ tmp, const float* __restrict__ residual,
                                                     CgState* st, RedCtx ctx) {
  MF_PDL_PROLOGUE
  __shared__ double sh[32];
  if (st->done) return;
  double acc[1] = { 0.0 };
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stp = (long long)gridDim.x * blockDim.x;
  if (V4) {
    for (const long long nq = n >> 2; i < nq; i += stp) {
      const float4 a = reinterpret_cast<const float4*>(tmp)[i], b = reinterpret_cast<const float4*>(residual)[i];
      acc[0] += (double)(a.x * b.x); acc[0] += (double)(a.y * b.y); acc[0] += (double)(a.z * b.z); acc[0] += (double)(a.w * b.w);
    }
  } else {
    for (; i < n; i += stp) acc[0] += (double)(tmp[i] * residual[i]);
  }
  const int ops[1] = { 0 };
  gridReduceK<1>(acc, ops, ctx, sh, FinBeta{ st });
}

// init: search = tmp (when given), sigma = tmp . residual                 (conjugategrad.cpp:229-234)
__global__ void __launch_bounds__(256) k_cg_dot_sigma(long long n, const float* __restrict__ tmp, const float* __restrict__ residual,
                                                      float* __restrict__ search, CgState* st, RedCtx ctx) {
  __shared__ double sh[32];
  double acc[1] = { 0.0 };
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stp = (long long)gridDim.x * blockDim.x;
  if (search) { for (; i < n; i += stp) { const float t = tmp[i]; search[i] = t; acc[0] += (double)(t * residual[i]); } }
  else { for (; i < n; i += stp) acc[0] += (double)(tmp[i] * residual[i]); }
  const int ops[1] = { 0 };
  gridReduceK<1>(acc, ops, ctx, sh, FinSigma{ st });
}

// search = tmp + beta * search                                            (UpdateSearchVec, conjugategrad.cpp:193)
template <bool V4>
__global__ void __launch_bounds__(256) k_cg_search(long long n, float* __restrict__ search, const float* __restrict__ tmp, const CgState* st) {
  MF_PDL_PROLOGUE
  if (st->done) return;
  const float beta = st->beta;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stp = (long long)gridDim.x * blockDim.x;
  if (V4) {
    for (const long long nq = n >> 2; i < nq; i += stp) {
      float4 se = reinterpret_cast<float4*>(search)[i];
      const float4 t = reinterpret_cast<const float4*>(tmp)[i];
      se.x = t.x + beta * se.x; se.y = t.y + beta * se.y; se.z = t.z + beta * se.z; se.w = t.w + beta * se.w;
      reinterpret_cast<float4*>(search)[i] = se;
    }
  } else {
    for (; i < n; i += stp) search[i] = tmp[i] + beta * search[i];
  }
}

// ---- red-black modified incomplete Cholesky: the parallel stand-in for the sequential MIC(0) --------
// (InitPreconditionModifiedIncompCholesky2 / ApplyPrecondition..., conjugategrad.cpp:66-97,135-159, restated for
//  a 2-colour ordering: red cells (i+j+k even) have no "lower" neighbours, black cells see all six red
//  neighbours as lower ones; tau = 0.97, sigma = 0.25 as in the reference.)
__device__ __forceinline__ float rbCoupling(const Geo& g, long long idx, int dir, const float* Ai, const float* Aj, const float* Ak, long long& n) {
  const long long X = 1, Y = g.nx, Z = g.Z;
  switch (dir) {
    case 0: n = idx - X; return Ai[n];
    case 1: n = idx + X; return Ai[idx];
    case 2: n = idx - Y; return Aj[n];
    case 3: n = idx + Y; return Aj[idx];
    case 4: n = idx - Z; return Ak[n];
    default: n = idx + Z; return Ak[idx];
  }
}
__device__ __forceinline__ float rbOffdiagSum(const Geo& g, long long n, const float* Ai, const float* Aj, const float* Ak) {
  const long long X = 1, Y = g.nx, Z = g.Z;
  return Ai[n - X] + Ai[n] + Aj[n - Y] + Aj[n] + Ak[n - Z] + Ak[n];
}

__global__ void __launch_bounds__(256) k_rbic_init(Geo g, const int* __restrict__ flags, float* __restrict__ P, const float* __restrict__ A0,
                                                   const float* __restrict__ Ai, const float* __restrict__ Aj, const float* __restrict__ Ak) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  float p = 0.f;
  if (interiorCell(g, i, j, k) && fIsFluid(flags[idx]) && A0[idx] > 0.f) {
    const float a0 = A0[idx];
    float e = a0;
    if ((i + j + k) & 1) {          // black: all neighbours are red, whose P = 1/sqrt(A0)
      const float tau = 0.97f, sigma = 0.25f;
      float sq = 0.f, fill = 0.f;
      for (int d = 0; d < 6; d++) {
        long long n; const float c = rbCoupling(g, idx, d, Ai, Aj, Ak, n);
        if (c != 0.f && A0[n] > 0.f) {
          const float pn2 = 1.0f / A0[n];
          sq += c * c * pn2;
          fill += c * (rbOffdiagSum(g, n, Ai, Aj, Ak) - c) * pn2;
        }
      }
      e = a0 - sq - tau * fill;
      if (e < sigma * a0) e = a0;
    }
    p = 1.0f / sqrtf(e);
  }
  P[idx] = p;
}

// forward substitution on black cells (red part is pointwise and folded in): q_b = P_b (r_b - sum_n A_bn P_n^2 r_n); z_b = P_b q_b
__global__ void __launch_bounds__(256) k_rbic_black(Geo g, const float* __restrict__ P, const float* __restrict__ r, float* __restrict__ z,
                                                    const float* __restrict__ Ai, const float* __restrict__ Aj, const float* __restrict__ Ak,
                                                    const CgState* st) {
  if (st->done) return;
  const int ii = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  const int i = 2 * ii + ((j + k + 1) & 1);       // black cells of this row
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  const float p = P[idx];
  if (p == 0.f) { z[idx] = 0.f; return; }
  float acc = r[idx];
  for (int d = 0; d < 6; d++) {
    long long n; const float c = rbCoupling(g, idx, d, Ai, Aj, Ak, n);
    if (c != 0.f) { const float pn = P[n]; acc -= c * pn * pn * r[n]; }
  }
  z[idx] = p * p * acc;
}
// backward substitution on red cells: z_r = P_r (P_r r_r - P_r sum_b A_rb z_b)
__global__ void __launch_bounds__(256) k_rbic_red(Geo g, const float* __restrict__ P, const float* __restrict__ r, float* __restrict__ z,
                                                  const float* __restrict__ Ai, const float* __restrict__ Aj, const float* __restrict__ Ak,
                                                  const CgState* st) {
  if (st->done) return;
  const int ii = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  const int i = 2 * ii + ((j + k) & 1);           // red cells of this row
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  const float p = P[idx];
  if (p == 0.f) { z[idx] = 0.f; return; }
  float acc = r[idx];
  for (int d = 0; d < 6; d++) {
    long long n; const float c = rbCoupling(g, idx, d, Ai, Aj, Ak, n);
    if (c != 0.f) acc -= c * z[n];
  }
  z[idx] = p * p * acc;
}

} // namespace mf
