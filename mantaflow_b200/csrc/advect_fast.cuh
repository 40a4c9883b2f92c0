// advect_fast.cuh -- the common configuration of advectSemiLagrange (3-D, orderSpace 1, orderTrace 1, fewer than 2^31 cells)
// with 32-bit index arithmetic and the clamp mode as a template parameter.  Same operations in the same order as the generic
// kernels of advect.cu (which stay for 2-D, cubic interpolation, second-order tracing and Vec3 grids) -- the advection kernels
// are instruction-issue bound, so what is saved here is address arithmetic and selects, not bytes:
//   * one int cell index; neighbours are immediate or single-IMAD offsets from per-plane base pointers,
//   * BUILD_INDEX (interpol.h:52-69) with s0 = 1 - s1 taken AFTER the edge overrides (1 - 0 = 1, 1 - 1 = 0: the same values the
//     reference assigns) and the index clamped with integer min/max (p < 0 <=> (int)p <= 0 is clamped to 0; p >= n-1 <=> (int)p >= n-1),
//   * clampMode 2 only compares against the min / max of the source cells, so they are taken with FMNMX (a NaN source value is
//     skipped by both forms; the sign of a zero bound cannot change a comparison); clampMode 1 returns the bounds and keeps the
//     reference's compare-and-assign form,
//   * the backward trace of MacCormack is p0 - v * (-dt) in the reference = p0 + v * dt bit for bit, so the face velocity times dt
//     is computed once for the trace and for the clamp.
// Reference: source/plugin/advection.cpp (SemiLagrange :25, SemiLagrangeMAC :45, MacCormackCorrect[MAC] :82/:95,
// doClampComponent[MAC] :146/:193, MacCormackClamp[MAC] :242/:271), source/util/interpol.h:52-94, source/grid.h:460-506.
#pragma once
#include "common.cuh"
#include <cfloat>

namespace mf {

struct GeoF { int nx, ny, nz, Z, k0, k1; };       // Z = nx * ny
__host__ __device__ inline GeoF geoF(const Geo& g) { GeoF f; f.nx = g.nx; f.ny = g.ny; f.nz = g.nz; f.Z = (int)g.Z; f.k0 = g.k0; f.k1 = g.k1; return f; }

#define FCHECKFLAG(f) (((f) & (MF_FLUID | MF_EMPTY)) != 0)

// one axis of BUILD_INDEX; p = pos - 0.5
__device__ __forceinline__ void axisW(float p, int n, int& xi, float& s0, float& s1) {
  const int t = (int)p;
  s1 = p - (float)t;
  if (p < 0.0f) s1 = 0.0f;
  if (p >= (float)(n - 1)) s1 = 1.0f;
  s0 = 1.0f - s1;
  xi = min(max(t, 0), n - 2);
}
// interpol() / interpolComponent<c>() on one plane (interpol.h:72-94), 3-D
__device__ __forceinline__ float interpolF(const float* __restrict__ d, const GeoF& g, float px, float py, float pz) {
  int xi, yi, zi; float s0, s1, t0, t1, f0, f1;
  axisW(px - 0.5f, g.nx, xi, s0, s1);
  axisW(py - 0.5f, g.ny, yi, t0, t1);
  axisW(pz - 0.5f, g.nz, zi, f0, f1);
  const float* p = d + (xi + g.nx * yi + g.Z * zi);
  const float* q = p + g.Z;
  const int Y = g.nx;
  return ((p[0] * t0 + p[Y] * t1) * s0 + (p[1] * t0 + p[Y + 1] * t1) * s1) * f0
       + ((q[0] * t0 + q[Y] * t1) * s0 + (q[1] * t0 + q[Y + 1] * t1) * s1) * f1;
}

#define IJK_FROM_BLOCK_F \
  MF_PDL_PROLOGUE \
  const GeoF g = geoF(gg); \
  const int i = blockIdx.x * blockDim.x + threadIdx.x; \
  const int j = blockIdx.y * blockDim.y + threadIdx.y; \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0; \
  if (i >= g.nx || j >= g.ny || k >= g.k1) return; \
  const int idx = i + g.nx * (j + g.ny * k); \
  const bool border = i < 1 || j < 1 || k < 1 || i >= g.nx - 1 || j >= g.ny - 1 || k >= g.nz - 1;

// ---- centred Real grids -------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_sl_f(Geo gg, CV3 vel, const float* __restrict__ src, float* __restrict__ dst, float dt) {
  IJK_FROM_BLOCK_F
  if (border) { dst[idx] = 0.0f; return; }
  const float* vx = vel.x + idx; const float* vy = vel.y + idx; const float* vz = vel.z + idx;
  const float cx = 0.5f * (vx[0] + vx[1]), cy = 0.5f * (vy[0] + vy[g.nx]), cz = 0.5f * (vz[0] + vz[g.Z]);     // getCentered
  dst[idx] = interpolF(src, g, (i + 0.5f) - cx * dt, (j + 0.5f) - cy * dt, (k + 0.5f) - cz * dt);
}

// MacCormackCorrect (advection.cpp:91,113): dst = fwd + strength * 0.5 * (old - bwd) -- `strength * 0.5` is a double, so the
// reference evaluates float(double(fw) + s2 * double(old - bwd)).  When s2 is a power of two (strength 1: s2 = 0.5) the float
// expression fw + s2f * t gives the same bits: s2f * t is exact, and the double sum is either exact (exponents at most 29 apart:
// 24 + 29 bits fit 53) or its smaller term is below 2^-29 of the larger one, far inside half a float ulp, so both roundings end at
// fw's neighbour-free value.  The only exception is a t so small that s2f * t underflows; that case keeps the double expression.
// 2 float operations instead of 3 conversions + 2 double operations per component (the FP64 / conversion pipes are quarter rate).
__device__ __forceinline__ float mcCorrect(float fw, float t, double s2, float s2f) {
  if (s2f != 0.f && !(fabsf(t) < 1.0e-30f && t != 0.f)) return fw + s2f * t;
  return (float)((double)fw + s2 * (double)t);
}
// strength * 0.5 as a float when it is a power of two (exact scaling), else 0 (host side; MF_ADVECT_MC_DOUBLE=1 keeps the double form)
inline float mcPow2Scale(float strength) {
  const double s2 = (double)strength * 0.5;
  const float f = (float)s2;
  int e = 0;
  const double m = frexp(s2, &e);
  return ((double)f == s2 && (m == 0.5 || m == -0.5) && e > -20 && e < 20 && !getenv("MF_ADVECT_MC_DOUBLE")) ? f : 0.f;
}

// backward SL + MacCormackCorrect + MacCormackClamp fused (see k_mc_centered in advect.cu)
template <int CM>
__global__ void __launch_bounds__(256) k_mc_centered_f(Geo gg, const int* __restrict__ flags, CV3 vel, const float* __restrict__ orig,
                                                       const float* __restrict__ fwd, float* __restrict__ dst, float dt, float strength, float s2f) {
  IJK_FROM_BLOCK_F
  const bool fluid = (flags[idx] & MF_FLUID) != 0;
  const double s2 = (double)strength * 0.5;
  if (border) {
    float d = 0.0f;
    if (fluid) d = (float)((double)0.0f + s2 * (double)(orig[idx] - 0.0f));
    dst[idx] = d;
    return;
  }
  const float* vxp = vel.x + idx; const float* vyp = vel.y + idx; const float* vzp = vel.z + idx;
  const float cx = 0.5f * (vxp[0] + vxp[1]), cy = 0.5f * (vyp[0] + vyp[g.nx]), cz = 0.5f * (vzp[0] + vzp[g.Z]);
  const float vx = cx * dt, vy = cy * dt, vz = cz * dt;
  const float fw = fwd[idx];
  float dv = fw;
  if (fluid) {
    const float bw = interpolF(fwd, g, (i + 0.5f) + vx, (j + 0.5f) + vy, (k + 0.5f) + vz);     // p0 - c * (-dt)
    dv = mcCorrect(fw, orig[idx] - bw, s2, s2f);
  }
  // doClampComponent (advection.cpp:146-187)
  float minv = FLT_MAX, maxv = -FLT_MAX;
  bool haveFl = false;
#pragma unroll
  for (int l = 0; l < (CM == 1 ? 2 : 1); l++) {
    const int cpx = (l == 0) ? (int)((float)i - vx) : (int)((float)i + vx);
    const int cpy = (l == 0) ? (int)((float)j - vy) : (int)((float)j + vy);
    const int cpz = (l == 0) ? (int)((float)k - vz) : (int)((float)k + vz);
    const int b = min(max(cpx, 0), g.nx - 2) + g.nx * min(max(cpy, 0), g.ny - 2) + g.Z * min(max(cpz, 0), g.nz - 2);
    const int* fp = flags + b; const float* op = orig + b;
#pragma unroll
    for (int dk = 0; dk < 2; dk++)
#pragma unroll
      for (int dj = 0; dj < 2; dj++)
#pragma unroll
        for (int di = 0; di < 2; di++) {
          const int o = di + g.nx * dj + g.Z * dk;
          if (FCHECKFLAG(fp[o])) {
            haveFl = true;
            const float v = op[o];
            if (CM == 1) { if (v < minv) minv = v; if (v > maxv) maxv = v; }
            else { minv = fminf(minv, v); maxv = fmaxf(maxv, v); }
          }
        }
  }
  if (!haveFl) dv = fw;
  else if (CM == 1) { if (dv < minv) dv = minv; else if (dv > maxv) dv = maxv; }
  else if ((dv < minv) | (dv > maxv)) dv = fw;
  if (CM == 1) {
    const int ux = g.nx - 1, uy = g.ny - 1, uz = g.nz - 1;
    const int fx = (int)(((float)i + 0.5f) - vx), fy = (int)(((float)j + 0.5f) - vy), fz = (int)(((float)k + 0.5f) - vz);
    const int bx = (int)(((float)i + 0.5f) + vx), by = (int)(((float)j + 0.5f) + vy), bz = (int)(((float)k + 0.5f) + vz);
    bool revert = fx < 0 || fy < 0 || fz < 0 || bx < 0 || by < 0 || bz < 0 || fx > ux || fy > uy || fz > uz || bx > ux || by > uy || bz > uz;
    if (!revert) revert = ((flags[fx + g.nx * fy + g.Z * fz] | flags[bx + g.nx * by + g.Z * bz]) & MF_OBSTACLE) != 0;
    if (revert) dv = fw;
  }
  dst[idx] = dv;
}

// ---- MAC grids ----------------------------------------------------------------------------------------------------
// getAtMACX/Y/Z (grid.h:473-506), 3-D, times dt
template <int C>
__device__ __forceinline__ void faceVelDt(const CV3& v, const GeoF& g, int idx, float dt, float& vx, float& vy, float& vz) {
  const float* x = v.x + idx; const float* y = v.y + idx; const float* z = v.z + idx;
  const int Y = g.nx, Z = g.Z;
  float cx, cy, cz;
  if (C == 0) {
    cx = x[0];
    cy = 0.25f * (y[0] + y[-1] + y[Y] + y[Y - 1]);
    cz = 0.25f * (z[0] + z[-1] + z[Z] + z[Z - 1]);
  } else if (C == 1) {
    cx = 0.25f * (x[0] + x[-Y] + x[1] + x[1 - Y]);
    cy = y[0];
    cz = 0.25f * (z[0] + z[-Y] + z[Z] + z[Z - Y]);
  } else {
    cx = 0.25f * (x[0] + x[-Z] + x[1] + x[1 - Z]);
    cy = 0.25f * (y[0] + y[-Z] + y[Y] + y[Y - Z]);
    cz = z[0];
  }
  vx = cx * dt; vy = cy * dt; vz = cz * dt;
}
__device__ __forceinline__ const float* compPtrF(const CV3& v, int c) { return c == 0 ? v.x : (c == 1 ? v.y : v.z); }

template <int C>
__device__ __forceinline__ float slMacCompF(const GeoF& g, const CV3& vel, const CV3& src, int i, int j, int k, int idx, float dt) {
  float vx, vy, vz;
  faceVelDt<C>(vel, g, idx, dt, vx, vy, vz);
  return interpolF(compPtrF(src, C), g, (i + 0.5f) - vx, (j + 0.5f) - vy, (k + 0.5f) - vz);
}
__global__ void __launch_bounds__(256) k_sl_mac_f(Geo gg, CV3 vel, CV3 src, V3 dst, float dt) {
  IJK_FROM_BLOCK_F
  if (border) { dst.x[idx] = 0.0f; dst.y[idx] = 0.0f; dst.z[idx] = 0.0f; return; }
  dst.x[idx] = slMacCompF<0>(g, vel, src, i, j, k, idx, dt);
  dst.y[idx] = slMacCompF<1>(g, vel, src, i, j, k, idx, dt);
  dst.z[idx] = slMacCompF<2>(g, vel, src, i, j, k, idx, dt);
}

// one component of backward SemiLagrangeMAC + MacCormackCorrectMAC + MacCormackClampMAC; fc / fn = flags of the cell and of its
// lower neighbour along the component's axis
template <int C, int CM>
__device__ __forceinline__ float mcMacCompF(const GeoF& g, const int* __restrict__ flags, const CV3& vel, const CV3& orig, const CV3& fwd,
                                            int i, int j, int k, int idx, float dt, double s2, float s2f, bool skip, int fc, int fn) {
  const float* fwdc = compPtrF(fwd, C);
  const float* origc = compPtrF(orig, C);
  const float fw = fwdc[idx];
  float vx, vy, vz;
  faceVelDt<C>(vel, g, idx, dt, vx, vy, vz);
  float d = fw;
  if (!skip) {
    const float bw = interpolF(fwdc, g, (i + 0.5f) + vx, (j + 0.5f) + vy, (k + 0.5f) + vz);     // p0 - v * (-dt)
    d = mcCorrect(fw, origc[idx] - bw, s2, s2f);                                               // advection.cpp:113
  }
  // doClampComponentMAC<C> (advection.cpp:193-236)
  if (CM == 2 && !(FCHECKFLAG(fc) && FCHECKFLAG(fn))) return fw;
  float minv = FLT_MAX, maxv = -FLT_MAX;
#pragma unroll
  for (int l = 0; l < (CM == 1 ? 2 : 1); l++) {
    const int cpx = (l == 0) ? (int)((float)i - vx) : (int)((float)i + vx);
    const int cpy = (l == 0) ? (int)((float)j - vy) : (int)((float)j + vy);
    const int cpz = (l == 0) ? (int)((float)k - vz) : (int)((float)k + vz);
    const float* op = origc + (min(max(cpx, 0), g.nx - 2) + g.nx * min(max(cpy, 0), g.ny - 2) + g.Z * min(max(cpz, 0), g.nz - 2));
    const float* oq = op + g.Z;
    const int Y = g.nx;
    const float o0 = op[0], o1 = op[1], o2 = op[Y], o3 = op[Y + 1], o4 = oq[0], o5 = oq[1], o6 = oq[Y], o7 = oq[Y + 1];
    if (CM == 1) {
      const float o[8] = { o0, o1, o2, o3, o4, o5, o6, o7 };
#pragma unroll
      for (int q = 0; q < 8; q++) { if (o[q] < minv) minv = o[q]; if (o[q] > maxv) maxv = o[q]; }
    } else {
      minv = fminf(fminf(fminf(o0, o1), fminf(o2, o3)), fminf(fminf(o4, o5), fminf(o6, o7)));
      maxv = fmaxf(fmaxf(fmaxf(o0, o1), fmaxf(o2, o3)), fmaxf(fmaxf(o4, o5), fmaxf(o6, o7)));
    }
  }
  if (CM == 1) { if (d < minv) return minv; if (d > maxv) return maxv; return d; }
  if (d < minv || d > maxv) return fw;
  return d;
}

template <int CM>
__global__ void __launch_bounds__(256, 8) k_mc_mac_f(Geo gg, const int* __restrict__ flags, CV3 vel, CV3 orig, CV3 fwd, V3 dst, float dt, float strength, float s2f) {
  IJK_FROM_BLOCK_F
  const int f = flags[idx];
  const int fxm = (i > 0) ? flags[idx - 1] : MF_FLUID, fym = (j > 0) ? flags[idx - g.nx] : MF_FLUID, fzm = (k > 0) ? flags[idx - g.Z] : MF_FLUID;
  const bool nf = (f & MF_FLUID) == 0;
  // skip[] of MacCormackCorrectMAC (advection.cpp:99-106)
  const bool sx = nf || (fxm & MF_FLUID) == 0, sy = nf || (fym & MF_FLUID) == 0, sz = nf || (fzm & MF_FLUID) == 0;
  const double s2 = (double)strength * 0.5;
  if (border) {
    // fwd == bwd == 0 on the border; MacCormackCorrectMAC still runs there (KERNEL() without bnd), the clamp does not
    dst.x[idx] = sx ? 0.0f : (float)((double)0.0f + s2 * (double)(orig.x[idx] - 0.0f));
    dst.y[idx] = sy ? 0.0f : (float)((double)0.0f + s2 * (double)(orig.y[idx] - 0.0f));
    dst.z[idx] = sz ? 0.0f : (float)((double)0.0f + s2 * (double)(orig.z[idx] - 0.0f));
    return;
  }
  dst.x[idx] = mcMacCompF<0, CM>(g, flags, vel, orig, fwd, i, j, k, idx, dt, s2, s2f, sx, f, fxm);
  dst.y[idx] = mcMacCompF<1, CM>(g, flags, vel, orig, fwd, i, j, k, idx, dt, s2, s2f, sy, f, fym);
  dst.z[idx] = mcMacCompF<2, CM>(g, flags, vel, orig, fwd, i, j, k, idx, dt, s2, s2f, sz, f, fzm);
}

} // namespace mf
