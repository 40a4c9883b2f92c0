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

// dst += alpha*search ; residual += -alpha*tmp ; norm(residual)     (conjugategrad.cpp:254-277)
// PCNONE: also tmp := residual (implicitly) and sigmaNew = residual . residual, beta        (:279-286)
template <bool PCNONE>
__global__ void __launch_bounds__(256) k_cg_update(Span sp, CgVecs v, CgState* st, RedCtx ctx) {
  MF_PDL_PROLOGUE
  __shared__ double sh[32];
  if (st->done) return;
  const float alpha = st->alpha, nalpha = -alpha;
  const int useL2 = st->useL2;
  double acc[4] = { 0.0, 1e300, -1e300, 0.0 };   // sumsqr, min, max, r.r (float products)
  const long long i0 = sp.off + blockIdx.x * (long long)blockDim.x + threadIdx.x, iend = sp.off + sp.n;
  const long long st_ = (long long)gridDim.x * blockDim.x;
  for (long long idx = i0; idx < iend; idx += st_) {
    v.dst[idx] += alpha * v.search[idx];
    const float r = v.residual[idx] + nalpha * v.tmp[idx];
    v.residual[idx] = r;
    if (useL2) { const double rd = (double)r; acc[0] += rd * rd; }
    else { const double rd = (double)r; if (rd < acc[1]) acc[1] = rd; if (rd > acc[2]) acc[2] = rd; }
    if (PCNONE) acc[3] += (double)(r * r);
  }
  const int ops[4] = { 0, 1, 2, 0 };
  gridReduceK<4>(acc, ops, ctx, sh, FinUpdate<PCNONE>{ st });
}

// sigmaNew = tmp . residual ; beta = sigmaNew / sigma                 (conjugategrad.cpp:279-295), preconditioned path
__global__ void __launch_bounds__(256) k_cg_dot_beta(long long n, const float* __restrict__ tmp, const float* __restrict__ residual,
                                                     CgState* st, RedCtx ctx) {
  MF_PDL_PROLOGUE
  __shared__ double sh[32];
  if (st->done) return;
  double acc[1] = { 0.0 };
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stp = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stp) acc[0] += (double)(tmp[i] * residual[i]);
  const int ops[1] = { 0 };
  gridReduceK<1>(acc, ops, ctx, sh, FinBeta{ st });
}

// init: sigma = tmp . residual                                           (conjugategrad.cpp:234)
__global__ void __launch_bounds__(256) k_cg_dot_sigma(long long n, const float* __restrict__ tmp, const float* __restrict__ residual,
                                                      CgState* st, RedCtx ctx) {
  __shared__ double sh[32];
  double acc[1] = { 0.0 };
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stp = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stp) acc[0] += (double)(tmp[i] * residual[i]);
  const int ops[1] = { 0 };
  gridReduceK<1>(acc, ops, ctx, sh, FinSigma{ st });
}

// search = tmp + beta * search                                            (UpdateSearchVec, conjugategrad.cpp:193)
__global__ void __launch_bounds__(256) k_cg_search(long long n, float* __restrict__ search, const float* __restrict__ tmp, const CgState* st) {
  MF_PDL_PROLOGUE
  if (st->done) return;
  const float beta = st->beta;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stp = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stp) search[i] = tmp[i] + beta * search[i];
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
