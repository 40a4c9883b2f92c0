// pressure_kernels.cuh -- streaming kernels of the pressure projection.
// Reference: source/plugin/pressure.cpp (MakeRhs :32, knCorrectVelocity :87, thetaHelper :115,
// ghostFluidHelper :123, surfTensHelper :130, ApplyGhostFluidDiagonal :136, knCorrectVelocityGhostFluid :154,
// knReplaceClampedGhostFluidVels :198, CountEmptyCells :217, fixPressure :226),
// source/conjugategrad.h (ApplyMatrix :118, ApplyMatrix2D :136, MakeLaplaceMatrix :154).
#pragma once
#include "common.cuh"
#include "reduce.cuh"
#include "quad.cuh"

namespace mf {

__device__ __forceinline__ bool interiorCell(const Geo& g, int i, int j, int k) {
  return i >= 1 && j >= 1 && i < g.nx - 1 && j < g.ny - 1 && (!g.is3D || (k >= 1 && k < g.nz - 1));
}

// ---- ghost fluid helpers (pressure.cpp:115-133) ------------------------------------------------------
__device__ __forceinline__ float thetaHelper(float inside, float outside) {
  const float denom = inside - outside;
  if ((double)denom > -1e-04) return 0.5f;
  return fmaxf(0.0f, fminf(1.0f, inside / denom));
}
__device__ __forceinline__ float ghostFluidHelper(const float* __restrict__ phi, long long idx, long long offset, float gfClamp) {
  const float alpha = thetaHelper(phi[idx], phi[idx + offset]);
  if (alpha < gfClamp) return gfClamp;
  return (float)(1. - (1. / (double)alpha));
}
__device__ __forceinline__ bool ghostFluidWasClamped(const float* __restrict__ phi, long long idx, long long offset, float gfClamp) {
  return thetaHelper(phi[idx], phi[idx + offset]) < gfClamp;
}
__device__ __forceinline__ float surfTensHelper(const float* __restrict__ phi, const float* __restrict__ curv, long long idx, long long offset,
                                                float surfTens, float gfClamp) {
  return surfTens * (curv[idx + offset] - ghostFluidHelper(phi, idx, offset, gfClamp) * curv[idx]);
}

struct RhsArgs {
  const int* flags; float* rhs; CV3 vel;
  const float* perCellCorr; CV3 fractions; int haveFractions; CV3 obvel; int haveObvel;
  const float* phi; const float* curv; float surfTens; float gfClamp;
  int writeBorder;     // 1: rhs is a fresh temp -> also write the zeros the reference's cleared grid holds on the border
};

// MakeRhs: bnd=1, reduce+ (double sum, int cnt)
__global__ void __launch_bounds__(256) k_make_rhs(Geo g, RhsArgs a, RedCtx ctx, double* out) {
  __shared__ double sh[32];
  double sum = 0.0, cnt = 0.0;
  MF_TILE_LOOP_BEGIN(g)
    if (!interiorCell(g, i, j, k)) { if (a.writeBorder) a.rhs[idx] = 0.0f; continue; }
    if (!fIsFluid(a.flags[idx])) { a.rhs[idx] = 0.0f; continue; }
    const long long X = 1, Y = g.nx, Z = g.Z;
    float set;
    if (!a.haveFractions) {
      set = a.vel.x[idx] - a.vel.x[idx + X] + a.vel.y[idx] - a.vel.y[idx + Y];
      if (g.is3D) set += a.vel.z[idx] - a.vel.z[idx + Z];
    } else {
      set = a.fractions.x[idx] * a.vel.x[idx] - a.fractions.x[idx + X] * a.vel.x[idx + X]
          + a.fractions.y[idx] * a.vel.y[idx] - a.fractions.y[idx + Y] * a.vel.y[idx + Y];
      if (g.is3D) set += a.fractions.z[idx] * a.vel.z[idx] - a.fractions.z[idx + Z] * a.vel.z[idx + Z];
      if (a.haveObvel) {
        set += (1 - a.fractions.x[idx]) * a.obvel.x[idx] - (1 - a.fractions.x[idx + X]) * a.obvel.x[idx + X]
             + (1 - a.fractions.y[idx]) * a.obvel.y[idx] - (1 - a.fractions.y[idx + Y]) * a.obvel.y[idx + Y];
        if (g.is3D) set += (1 - a.fractions.z[idx]) * a.obvel.z[idx] - (1 - a.fractions.z[idx + Z]) * a.obvel.z[idx + Z];
      }
    }
    if (a.phi && a.curv) {
      if (fIsEmpty(a.flags[idx - X])) set += surfTensHelper(a.phi, a.curv, idx, -X, a.surfTens, a.gfClamp);
      if (fIsEmpty(a.flags[idx + X])) set += surfTensHelper(a.phi, a.curv, idx, +X, a.surfTens, a.gfClamp);
      if (fIsEmpty(a.flags[idx - Y])) set += surfTensHelper(a.phi, a.curv, idx, -Y, a.surfTens, a.gfClamp);
      if (fIsEmpty(a.flags[idx + Y])) set += surfTensHelper(a.phi, a.curv, idx, +Y, a.surfTens, a.gfClamp);
      if (g.is3D) {
        if (fIsEmpty(a.flags[idx - Z])) set += surfTensHelper(a.phi, a.curv, idx, -Z, a.surfTens, a.gfClamp);
        if (fIsEmpty(a.flags[idx + Z])) set += surfTensHelper(a.phi, a.curv, idx, +Z, a.surfTens, a.gfClamp);
      }
    }
    if (a.perCellCorr) set += a.perCellCorr[idx];
    sum += (double)set;
    cnt += 1.0;
    a.rhs[idx] = set;
  MF_TILE_LOOP_END
  double acc[2] = { sum, cnt };
  const int ops[2] = { 0, 0 };
  gridReduceK<2>(acc, ops, ctx, sh, FinStore<2>{ out });
}

// MakeRhs of the plain projection (no fractions, obstacle velocities, surface tension or per-cell correction):
// W x-consecutive cells per thread (quad.cuh), same expression per cell
template <int W>
__global__ void __launch_bounds__(256) k_make_rhs_plain(Geo g, const int* __restrict__ flags, float* __restrict__ rhs, CV3 vel, int writeBorder,
                                                        RedCtx ctx, double* out) {
  MF_PDL_PROLOGUE
  __shared__ double sh[32];
  double sum = 0.0, cnt = 0.0;
  MF_QTILE_LOOP_BEGIN(g, W)
    const long long Y = g.nx, Z = g.Z;
    float r[W];
    if (j < 1 || j >= g.ny - 1 || (g.is3D && (k < 1 || k >= g.nz - 1))) {       // border row
      if (writeBorder && on) { for (int c = 0; c < W; c++) r[c] = 0.f; stW<W>(rhs, idx, r); }
      continue;
    }
    int f[W]; float vx[W], vxp[W], vy[W], vyp[W], vz[W], vzp[W];
    ldW<W>(f, flags, idx);
    ldW<W>(vx, vel.x, idx);
    ldW<W>(vy, vel.y, idx); ldW<W>(vyp, vel.y, idx + Y);
    if (g.is3D) { ldW<W>(vz, vel.z, idx); ldW<W>(vzp, vel.z, idx + Z); }
    // (the shuffle of the x + 1 neighbour behind all loads, see quad.cuh)
    if constexpr (W > 1) { const float re = xEdgePlus<W, float>(vel.x, idx, i, g.nx, 0.f); xPlusWS<W, float>(vxp, vx, re, i, g.nx); }
    else xPlusW<W, float>(vxp, vx, vel.x, idx, i, g.nx, 0.f);
    if (!on) continue;
    const bool edgeQuad = (i == 0) || (i + W >= g.nx);
    if (edgeQuad && !writeBorder) ldW<W>(r, rhs, idx);
#pragma unroll
    for (int c = 0; c < W; c++) {
      const int ic = i + c;
      if (ic < 1 || ic >= g.nx - 1) { if (writeBorder) r[c] = 0.f; continue; }
      if (!fIsFluid(f[c])) { r[c] = 0.f; continue; }
      float set = vx[c] - vxp[c] + vy[c] - vyp[c];
      if (g.is3D) set += vz[c] - vzp[c];
      sum += (double)set;
      cnt += 1.0;
      r[c] = set;
    }
    stW<W>(rhs, idx, r);
  MF_QTILE_LOOP_END
  double acc[2] = { sum, cnt };
  const int ops[2] = { 0, 0 };
  gridReduceK<2>(acc, ops, ctx, sh, FinStore<2>{ out });
}

__global__ void k_add_scalar(float* d, float v, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) d[i] += v;
}

// 16-bit row code of the plain (no fractions, no ghost fluid) Laplace matrix: every coefficient is 0 or -1 and the diagonal a
// small integer, so a row of ApplyMatrix is fully described by 10 bits.  The CG then streams 2 B/cell of matrix instead of the
// four float planes (16 B/cell) + flags (4 B/cell).  Bits: coefficient towards +x,+y,+z (Ai/Aj/Ak at the cell), towards
// -x,-y,-z (Ai/Aj/Ak at the lower neighbour), the diagonal (0..6), and "fluid" (non-fluid rows are identity rows, conjugategrad.h:217).
enum { CGC_XP = 1, CGC_YP = 2, CGC_ZP = 4, CGC_XM = 8, CGC_YM = 16, CGC_ZM = 32, CGC_DIAG_SHIFT = 6, CGC_FLUID = 512 };

// MakeLaplaceMatrix (+ ApplyGhostFluidDiagonal) -- writes all four coefficient grids for every cell
// (the reference accumulates into cleared temp grids, conjugategrad.h:154-187, pressure.cpp:136-151)
// CODE: also write the 16-bit row code (only valid without fractions / phi).
template <bool CODE>
__global__ void __launch_bounds__(256) k_make_laplace(Geo g, const int* __restrict__ flags, float* __restrict__ A0, float* __restrict__ Ai,
                                                      float* __restrict__ Aj, float* __restrict__ Ak, CV3 fr, int haveFractions,
                                                      const float* __restrict__ phi, float gfClamp, unsigned short* __restrict__ code) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  float a0 = 0.f, ai = 0.f, aj = 0.f, ak = 0.f;
  if (interiorCell(g, i, j, k) && fIsFluid(flags[idx])) {
    const long long X = 1, Y = g.nx, Z = g.Z;
    const int fxm = flags[idx - X], fxp = flags[idx + X], fym = flags[idx - Y], fyp = flags[idx + Y];
    const int fzm = g.is3D ? flags[idx - Z] : 0, fzp = g.is3D ? flags[idx + Z] : 0;
    if (!haveFractions) {
      if (!fIsObstacle(fxm)) a0 += 1.f;
      if (!fIsObstacle(fxp)) a0 += 1.f;
      if (!fIsObstacle(fym)) a0 += 1.f;
      if (!fIsObstacle(fyp)) a0 += 1.f;
      if (g.is3D && !fIsObstacle(fzm)) a0 += 1.f;
      if (g.is3D && !fIsObstacle(fzp)) a0 += 1.f;
      if (fIsFluid(fxp)) ai = -1.f;
      if (fIsFluid(fyp)) aj = -1.f;
      if (g.is3D && fIsFluid(fzp)) ak = -1.f;
    } else {
      a0 += fr.x[idx]; a0 += fr.x[idx + X];
      a0 += fr.y[idx]; a0 += fr.y[idx + Y];
      if (g.is3D) { a0 += fr.z[idx]; a0 += fr.z[idx + Z]; }
      if (fIsFluid(fxp)) ai = -fr.x[idx + X];
      if (fIsFluid(fyp)) aj = -fr.y[idx + Y];
      if (g.is3D && fIsFluid(fzp)) ak = -fr.z[idx + Z];
    }
    if (phi) {
      if (fIsEmpty(fxm)) a0 -= ghostFluidHelper(phi, idx, -X, gfClamp);
      if (fIsEmpty(fxp)) a0 -= ghostFluidHelper(phi, idx, +X, gfClamp);
      if (fIsEmpty(fym)) a0 -= ghostFluidHelper(phi, idx, -Y, gfClamp);
      if (fIsEmpty(fyp)) a0 -= ghostFluidHelper(phi, idx, +Y, gfClamp);
      if (g.is3D) {
        if (fIsEmpty(fzm)) a0 -= ghostFluidHelper(phi, idx, -Z, gfClamp);
        if (fIsEmpty(fzp)) a0 -= ghostFluidHelper(phi, idx, +Z, gfClamp);
      }
    }
  }
  A0[idx] = a0; Ai[idx] = ai; Aj[idx] = aj; Ak[idx] = ak;
  if (CODE) {
    // the -x/-y/-z coefficients live in the lower neighbour's row: Ai[idx-X] = -1 iff that cell is an interior fluid cell and this
    // one is fluid.  Neighbours outside the coordinate range map (by linear index) to border cells, whose rows are all zero.
    int c = 0;
    const int f = flags[idx];
    if (fIsFluid(f)) {
      c = CGC_FLUID | ((int)a0 << CGC_DIAG_SHIFT);
      if (ai != 0.f) c |= CGC_XP;
      if (aj != 0.f) c |= CGC_YP;
      if (ak != 0.f) c |= CGC_ZP;
      const long long Y = g.nx, Z = g.Z;
      if (i >= 1 && interiorCell(g, i - 1, j, k) && fIsFluid(flags[idx - 1])) c |= CGC_XM;
      if (j >= 1 && interiorCell(g, i, j - 1, k) && fIsFluid(flags[idx - Y])) c |= CGC_YM;
      if (g.is3D && k >= 1 && interiorCell(g, i, j, k - 1) && fIsFluid(flags[idx - Z])) c |= CGC_ZM;
    }
    code[idx] = (unsigned short)c;
  }
}

// Row codes only (no coefficient planes), W x-consecutive cells per thread: the matrix setup of a plain projection step whose
// preconditioner is already built.  Same bits as k_make_laplace<false, true>.
// the row codes of W x-consecutive cells; f = their flags (also returned for the caller's own use).  All lanes of a warp take part
// (shuffles), `on` = the quad lies inside the row
template <int W>
__device__ __forceinline__ void laplaceCodeQuad(const Geo& g, const int* __restrict__ flags, long long idx, int i, int j, int k, int (&f)[W], int (&out)[W]) {
  const long long Y = g.nx, Z = g.Z;
  int fxm[W], fxp[W], fym[W], fyp[W], fzm[W], fzp[W];
  ldW<W>(f, flags, idx);
  if (j > 0) ldW<W>(fym, flags, idx - Y); else { for (int c = 0; c < W; c++) fym[c] = 0; }
  if (j < g.ny - 1) ldW<W>(fyp, flags, idx + Y); else { for (int c = 0; c < W; c++) fyp[c] = 0; }
  if (g.is3D && k > 0) ldW<W>(fzm, flags, idx - Z); else { for (int c = 0; c < W; c++) fzm[c] = 0; }
  if (g.is3D && k < g.nz - 1) ldW<W>(fzp, flags, idx + Z); else { for (int c = 0; c < W; c++) fzp[c] = 0; }
  // (the shuffles of the x neighbours behind all loads, see quad.cuh)
  if constexpr (W > 1) {
    const int le = xEdgeMinus<W, int>(flags, idx, i, 0), re = xEdgePlus<W, int>(flags, idx, i, g.nx, 0);
    xMinusWS<W, int>(fxm, f, le); xPlusWS<W, int>(fxp, f, re, i, g.nx);
  } else { xMinusW<W, int>(fxm, f, flags, idx, i, 0); xPlusW<W, int>(fxp, f, flags, idx, i, g.nx, 0); }
  const bool rowJK = j >= 1 && j < g.ny - 1 && (!g.is3D || (k >= 1 && k < g.nz - 1));
#pragma unroll
  for (int c = 0; c < W; c++) {
    int cd = 0;
    if (fIsFluid(f[c])) {
      cd = CGC_FLUID;
      const int ic = i + c;
      if (rowJK && ic >= 1 && ic < g.nx - 1) {          // an interior fluid cell has a Laplace row
        int a0 = 0;
        a0 += !fIsObstacle(fxm[c]); a0 += !fIsObstacle(fxp[c]); a0 += !fIsObstacle(fym[c]); a0 += !fIsObstacle(fyp[c]);
        if (g.is3D) { a0 += !fIsObstacle(fzm[c]); a0 += !fIsObstacle(fzp[c]); }
        cd |= a0 << CGC_DIAG_SHIFT;
        if (fIsFluid(fxp[c])) cd |= CGC_XP;
        if (fIsFluid(fyp[c])) cd |= CGC_YP;
        if (g.is3D && fIsFluid(fzp[c])) cd |= CGC_ZP;
      }
      // couplings stored in the lower neighbours' rows (those cells must be interior fluid cells)
      if (rowJK && ic >= 2 && ic < g.nx && fIsFluid(fxm[c])) cd |= CGC_XM;                         // (ic-1, j, k) interior
      if (ic >= 1 && ic < g.nx - 1 && j >= 2 && (!g.is3D || (k >= 1 && k < g.nz - 1)) && fIsFluid(fym[c])) cd |= CGC_YM;   // (ic, j-1, k) interior
      if (g.is3D && ic >= 1 && ic < g.nx - 1 && j >= 1 && j < g.ny - 1 && k >= 2 && fIsFluid(fzm[c])) cd |= CGC_ZM;       // (ic, j, k-1) interior
    }
    out[c] = cd;
  }
}
template <int W>
__global__ void __launch_bounds__(256) k_make_laplace_code(Geo g, const int* __restrict__ flags, unsigned short* __restrict__ code) {
  MF_CELLS_W(g, W)
  int f[W], out[W];
  laplaceCodeQuad<W>(g, flags, idx, i, j, k, f, out);
  if (!on) return;
  stW<W>(code, idx, out);
}
// The same in the pass that gathers the statistics zeroPressureFixing needs (k_flag_stats below: CountEmptyCells + the first
// interior fluid cell): one read of the flag grid instead of two.  1-D grid-stride tile loop (the reduction's partials are bounded).
template <int W>
__global__ void __launch_bounds__(256) k_make_laplace_code_stats(Geo g, const int* __restrict__ flags, unsigned short* __restrict__ code, RedCtx ctx, double* outStats) {
  __shared__ double sh[32];
  double numEmpty = 0.0, minFluid = 1e300;
  MF_QTILE_LOOP_BEGIN(g, W)
    int f[W], out[W];
    laplaceCodeQuad<W>(g, flags, idx, i, j, k, f, out);
    if (!on) continue;
    stW<W>(code, idx, out);
    const bool rowJK = j >= 1 && j < g.ny - 1 && (!g.is3D || (k >= 1 && k < g.nz - 1));
#pragma unroll
    for (int c = 0; c < W; c++) {
      if (fIsEmpty(f[c])) numEmpty += 1.0;
      if (fIsFluid(f[c]) && rowJK && i + c >= 1 && i + c < g.nx - 1) { const double v = (double)(idx + c); if (v < minFluid) minFluid = v; }
    }
  MF_QTILE_LOOP_END
  double acc[2] = { numEmpty, minFluid };
  const int ops[2] = { 0, 1 };
  gridReduceK<2>(acc, ops, ctx, sh, FinStore<2>{ outStats });
}

// CountEmptyCells (pressure.cpp:217) + "first interior fluid cell in FOR_IJK_BND order" (:371-381)
// out[0] = numEmpty, out[1] = min interior fluid index (or INT_MAX-ish)
template <int W>
__global__ void __launch_bounds__(256) k_flag_stats(Geo g, const int* __restrict__ flags, RedCtx ctx, double* out) {
  __shared__ double sh[32];
  double numEmpty = 0.0, minFluid = 1e300;
  MF_QTILE_LOOP_BEGIN(g, W)
    if (!on) continue;
    int f[W];
    ldW<W>(f, flags, idx);
    const bool rowJK = j >= 1 && j < g.ny - 1 && (!g.is3D || (k >= 1 && k < g.nz - 1));
#pragma unroll
    for (int c = 0; c < W; c++) {
      if (fIsEmpty(f[c])) numEmpty += 1.0;
      if (fIsFluid(f[c]) && rowJK && i + c >= 1 && i + c < g.nx - 1) { const double v = (double)(idx + c); if (v < minFluid) minFluid = v; }
    }
  MF_QTILE_LOOP_END
  double acc[2] = { numEmpty, minFluid };
  const int ops[2] = { 0, 1 };
  gridReduceK<2>(acc, ops, ctx, sh, FinStore<2>{ out });
}

// zeroPressureFixing: pick the cell (pressure.cpp:352-383) and apply fixPressure(idx, 0, ...) (:226-245); one thread
template <bool PLANES, bool CODE>
__global__ void k_fix_pressure(Geo g, const int* __restrict__ flags, const double* __restrict__ stats, float* rhs, float* A0, float* Ai,
                               float* Aj, float* Ak, unsigned short* code, int* fixedOut) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  long long fix = -1;
  if (stats[0] == 0.0) {
    const int tx = g.nx / 2, ty = g.ny - 1, tz = g.is3D ? g.nz / 2 : 0;
    for (int q = 0; q < 3 && fix < 0; q++) {
      const int y = ty - q;
      if (y < 0) continue;
      const long long p = (long long)tx + (long long)g.nx * y + g.Z * tz;
      if (fIsFluid(flags[p])) fix = p;
    }
    if (fix < 0 && stats[1] < 1e299) fix = (long long)stats[1];
  }
  *fixedOut = (int)fix;
  if (fix < 0) return;
  const float value = 0.0f;
  const long long X = 1, Y = g.nx, Z = g.Z;
  if (PLANES) {
    rhs[fix + X] -= Ai[fix] * value;
    rhs[fix + Y] -= Aj[fix] * value;
    rhs[fix - X] -= Ai[fix - X] * value;
    rhs[fix - Y] -= Aj[fix - Y] * value;
    if (g.is3D) { rhs[fix + Z] -= Ak[fix] * value; rhs[fix - Z] -= Ak[fix - Z] * value; }
    A0[fix] = 1.0f;
    Ai[fix] = Aj[fix] = Ak[fix] = 0.0f;
    Ai[fix - X] = 0.0f;
    Aj[fix - Y] = 0.0f;
    if (g.is3D) Ak[fix - Z] = 0.0f;
  }
  rhs[fix] = value;      // (without the planes the neighbour updates above are rhs -= (0 or -1) * 0, i.e. no-ops)
  if (CODE) {
    // same edits on the row codes: the row becomes (1, 0, 0, 0) and all six couplings to the cell are cut
    code[fix] = (unsigned short)((code[fix] & CGC_FLUID) | (1 << CGC_DIAG_SHIFT));
    code[fix + X] &= (unsigned short)~CGC_XM; code[fix - X] &= (unsigned short)~CGC_XP;
    code[fix + Y] &= (unsigned short)~CGC_YM; code[fix - Y] &= (unsigned short)~CGC_YP;
    if (g.is3D) { code[fix + Z] &= (unsigned short)~CGC_ZM; code[fix - Z] &= (unsigned short)~CGC_ZP; }
  }
}

// ApplyMatrix / ApplyMatrix2D for one cell; neighbours outside the array contribute nothing
// (the reference reads out of bounds there, which only happens for fluid cells on the domain border)
__device__ __forceinline__ float applyRow(const Geo& g, long long idx, const float* __restrict__ src, const float* __restrict__ A0,
                                          const float* __restrict__ Ai, const float* __restrict__ Aj, const float* __restrict__ Ak) {
  const long long X = 1, Y = g.nx, Z = g.Z, N = g.N;
  float r = src[idx] * A0[idx];
  r += (idx - X >= 0) ? src[idx - X] * Ai[idx - X] : 0.0f;
  r += (idx + X < N) ? src[idx + X] * Ai[idx] : 0.0f;
  r += (idx - Y >= 0) ? src[idx - Y] * Aj[idx - Y] : 0.0f;
  r += (idx + Y < N) ? src[idx + Y] * Aj[idx] : 0.0f;
  if (g.is3D) {
    r += (idx - Z >= 0) ? src[idx - Z] * Ak[idx - Z] : 0.0f;
    r += (idx + Z < N) ? src[idx + Z] * Ak[idx] : 0.0f;
  }
  return r;
}

__global__ void __launch_bounds__(256) k_apply_matrix(Geo g, const int* __restrict__ flags, float* __restrict__ dst, const float* __restrict__ src,
                                                      const float* __restrict__ A0, const float* __restrict__ Ai, const float* __restrict__ Aj,
                                                      const float* __restrict__ Ak) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  dst[idx] = fIsFluid(flags[idx]) ? applyRow(g, idx, src, A0, Ai, Aj, Ak) : src[idx];
}

// knCorrectVelocity (pressure.cpp:87-109), in place; W x-consecutive cells per thread (quad.cuh)
template <int W>
__global__ void __launch_bounds__(256) k_correct_velocity(Geo g, const int* __restrict__ flags, V3 vel, const float* __restrict__ p) {
  MF_PDL_PROLOGUE
  MF_CELLS_W(g, W)
  if (j < 1 || j >= g.ny - 1 || (g.is3D && (k < 1 || k >= g.nz - 1))) return;     // bnd = 1 rows only (warp-uniform)
  const long long Y = g.nx, Z = g.Z;
  // all loads first, the shuffles of the x - 1 neighbours behind them (see quad.cuh)
  int f[W], fx[W], fy[W], fz[W];
  float pc[W], px[W], py[W], pz[W];
  float vx[W], vy[W], vz[W];
  ldW<W>(f, flags, idx); ldW<W>(fy, flags, idx - Y);
  if (g.is3D) ldW<W>(fz, flags, idx - Z);
  ldW<W>(pc, p, idx); ldW<W>(py, p, idx - Y);
  if (g.is3D) ldW<W>(pz, p, idx - Z);
  ldW<W>(vx, vel.x, idx); ldW<W>(vy, vel.y, idx);
  if (g.is3D) ldW<W>(vz, vel.z, idx);
  if constexpr (W > 1) {
    const int fe = xEdgeMinus<W, int>(flags, idx, i, 0);
    const float pe = xEdgeMinus<W, float>(p, idx, i, 0.f);
    xMinusWS<W, int>(fx, f, fe); xMinusWS<W, float>(px, pc, pe);
  } else { xMinusW<W, int>(fx, f, flags, idx, i, 0); xMinusW<W, float>(px, pc, p, idx, i, 0.f); }
  if (!g.is3D) { for (int c = 0; c < W; c++) { fz[c] = 0; pz[c] = 0.f; } }
#pragma unroll
  for (int c = 0; c < W; c++) {
    if (i + c < 1 || i + c >= g.nx - 1) continue;
    if (fIsFluid(f[c])) {
      if (fIsFluid(fx[c])) vx[c] -= (pc[c] - px[c]);
      if (fIsFluid(fy[c])) vy[c] -= (pc[c] - py[c]);
      if (fIsEmpty(fx[c])) vx[c] -= pc[c];
      if (fIsEmpty(fy[c])) vy[c] -= pc[c];
      if (g.is3D) {
        if (fIsFluid(fz[c])) vz[c] -= (pc[c] - pz[c]);
        if (fIsEmpty(fz[c])) vz[c] -= pc[c];
      }
    } else if (fIsEmpty(f[c]) && !fIsOutflow(f[c])) {
      if (fIsFluid(fx[c])) vx[c] += px[c]; else vx[c] = 0.f;
      if (fIsFluid(fy[c])) vy[c] += py[c]; else vy[c] = 0.f;
      if (g.is3D) { if (fIsFluid(fz[c])) vz[c] += pz[c]; else vz[c] = 0.f; }
    }
  }
  if (!on) return;
  stW<W>(vel.x, idx, vx); stW<W>(vel.y, idx, vy);
  if (g.is3D) stW<W>(vel.z, idx, vz);
}

// knCorrectVelocityGhostFluid (pressure.cpp:154-187)
__global__ void __launch_bounds__(256) k_correct_velocity_gf(Geo g, const int* __restrict__ flags, V3 vel, const float* __restrict__ p,
                                                             const float* __restrict__ phi, float gfClamp, const float* __restrict__ curv, float surfTens) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  if (!interiorCell(g, i, j, k)) return;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  const long long X = 1, Y = g.nx, Z = g.Z;
  const int f = flags[idx];
  const int fx = flags[idx - X], fy = flags[idx - Y], fz = g.is3D ? flags[idx - Z] : 0;
  const bool fluid = fIsFluid(f), emptyNoOut = fIsEmpty(f) && !fIsOutflow(f);
  if (fluid) {
    if (fIsEmpty(fx)) vel.x[idx] += p[idx] * ghostFluidHelper(phi, idx, -X, gfClamp);
    if (fIsEmpty(fy)) vel.y[idx] += p[idx] * ghostFluidHelper(phi, idx, -Y, gfClamp);
    if (g.is3D && fIsEmpty(fz)) vel.z[idx] += p[idx] * ghostFluidHelper(phi, idx, -Z, gfClamp);
  } else if (emptyNoOut) {
    if (fIsFluid(fx)) vel.x[idx] -= p[idx - X] * ghostFluidHelper(phi, idx - X, +X, gfClamp); else vel.x[idx] = 0.f;
    if (fIsFluid(fy)) vel.y[idx] -= p[idx - Y] * ghostFluidHelper(phi, idx - Y, +Y, gfClamp); else vel.y[idx] = 0.f;
    if (g.is3D) { if (fIsFluid(fz)) vel.z[idx] -= p[idx - Z] * ghostFluidHelper(phi, idx - Z, +Z, gfClamp); else vel.z[idx] = 0.f; }
  }
  if (curv) {
    if (fluid) {
      if (fIsEmpty(fx)) vel.x[idx] += surfTensHelper(phi, curv, idx, -X, surfTens, gfClamp);
      if (fIsEmpty(fy)) vel.y[idx] += surfTensHelper(phi, curv, idx, -Y, surfTens, gfClamp);
      if (g.is3D && fIsEmpty(fz)) vel.z[idx] += surfTensHelper(phi, curv, idx, -Z, surfTens, gfClamp);
    } else if (emptyNoOut) {
      vel.x[idx] -= fIsFluid(fx) ? surfTensHelper(phi, curv, idx - X, +X, surfTens, gfClamp) : 0.f;
      vel.y[idx] -= fIsFluid(fy) ? surfTensHelper(phi, curv, idx - Y, +Y, surfTens, gfClamp) : 0.f;
      if (g.is3D) vel.z[idx] -= fIsFluid(fz) ? surfTensHelper(phi, curv, idx - Z, +Z, surfTens, gfClamp) : 0.f;
    }
  }
}

// knReplaceClampedGhostFluidVels (pressure.cpp:198-214): empty cells copy face velocities from fluid neighbours
__global__ void __launch_bounds__(256) k_replace_clamped_gf(Geo g, const int* __restrict__ flags, V3 vel, const float* __restrict__ phi, float gfClamp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  if (!interiorCell(g, i, j, k)) return;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  const long long X = 1, Y = g.nx, Z = g.Z;
  if (!fIsEmpty(flags[idx])) return;
  if (fIsFluid(flags[idx - X]) && ghostFluidWasClamped(phi, idx - X, +X, gfClamp)) vel.x[idx] = vel.x[idx - X];
  if (fIsFluid(flags[idx - Y]) && ghostFluidWasClamped(phi, idx - Y, +Y, gfClamp)) vel.y[idx] = vel.y[idx - Y];
  if (g.is3D && fIsFluid(flags[idx - Z]) && ghostFluidWasClamped(phi, idx - Z, +Z, gfClamp)) vel.z[idx] = vel.z[idx - Z];
  if (fIsFluid(flags[idx + X]) && ghostFluidWasClamped(phi, idx + X, -X, gfClamp)) vel.x[idx] = vel.x[idx + X];
  if (fIsFluid(flags[idx + Y]) && ghostFluidWasClamped(phi, idx + Y, -Y, gfClamp)) vel.y[idx] = vel.y[idx + Y];
  if (g.is3D && fIsFluid(flags[idx + Z]) && ghostFluidWasClamped(phi, idx + Z, -Z, gfClamp)) vel.z[idx] = vel.z[idx + Z];
}

} // namespace mf
