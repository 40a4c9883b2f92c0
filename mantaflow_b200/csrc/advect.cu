// advect.cu -- semi-Lagrangian and MacCormack advection (reference: source/plugin/advection.cpp).
//
// Kernel plan (DESIGN.md "advection"):
//   order 1 : one kernel, dst = SL(src)                          (SemiLagrange :25 / SemiLagrangeMAC :45)
//   order 2 : K1 fwd = SL(orig, +dt); K2 fuses the reference's three passes
//             bwd = SL(fwd, -dt) (:312/:426), MacCormackCorrect[MAC] (:82/:95) and
//             MacCormackClamp[MAC] (:242/:271) per cell -- bwd and the un-clamped value never touch HBM.
//   Both kernels write every cell (border cells get the value the reference's cleared temp grid
//   would hold), so no memset of temporaries is needed.
//   MAC grids: applyOutflowBC (:388) as two kernels, launched only when the flag grid holds outflow cells
//   (cached per flag grid, invalidated by every flag-writing entry point).
#include "common.cuh"
#include "sample.cuh"
#include "advect_fast.cuh"
#include "advect_tma.cuh"
#include <cfloat>

namespace mf {

template <int OS, bool VEC>
__device__ __forceinline__ float sampleHi(const float* __restrict__ d, const Geo& g, float x, float y, float z) {
  if (OS == 1) return interpol(d, g, x, y, z);
  return interpolCubic<VEC>(d, g, x, y, z);
}

__device__ __forceinline__ bool isBorder(const Geo& g, int i, int j, int k) {
  return i < 1 || j < 1 || i >= g.nx - 1 || j >= g.ny - 1 || (g.is3D && (k < 1 || k >= g.nz - 1));
}

#define IJK_FROM_BLOCK \
  const int i = blockIdx.x * blockDim.x + threadIdx.x; \
  const int j = blockIdx.y * blockDim.y + threadIdx.y; \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0; \
  if (i >= g.nx || j >= g.ny || k >= g.k1) return; \
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;

// back-traced position for centred grids (advection.cpp:28-38)
template <int OT>
__device__ __forceinline__ void tracePos(const Geo& g, const CV3& vel, int i, int j, int k, long long idx, float dt,
                                         float& px, float& py, float& pz) {
  float cx, cy, cz;
  getCentered(vel, g, idx, cx, cy, cz);
  const float p0x = i + 0.5f, p0y = j + 0.5f, p0z = k + 0.5f;
  if (OT == 1) {
    px = p0x - cx * dt; py = p0y - cy * dt; pz = p0z - cz * dt;
  } else {
    const float p1x = p0x - (cx * dt) * 0.5f, p1y = p0y - (cy * dt) * 0.5f, p1z = p0z - (cz * dt) * 0.5f;
    float vx, vy, vz;
    interpolMAC(vel.x, vel.y, vel.z, g, p1x, p1y, p1z, vx, vy, vz);
    px = p0x - vx * dt; py = p0y - vy * dt; pz = p0z - vz * dt;
  }
}

// ---- SemiLagrange<T>, T = Real (NC=1) or Vec3 (NC=3) --------------------------------------
template <int NC, int OS, int OT>
__global__ void __launch_bounds__(256) k_sl(Geo g, CV3 vel, const float* __restrict__ src, float* __restrict__ dst, float dt, long long np) {
  IJK_FROM_BLOCK
  if (isBorder(g, i, j, k)) {
#pragma unroll
    for (int c = 0; c < NC; c++) dst[idx + c * np] = 0.0f;
    return;
  }
  float px, py, pz;
  tracePos<OT>(g, vel, i, j, k, idx, dt, px, py, pz);
#pragma unroll
  for (int c = 0; c < NC; c++) dst[idx + c * np] = sampleHi<OS, (NC == 3)>(src + c * np, g, px, py, pz);
}

#define CHECKFLAG(f) (((f) & (MF_FLUID | MF_EMPTY)) != 0)

// fused backward SL + MacCormackCorrect + MacCormackClamp for centred grids
template <int NC, int OS, int OT>
__global__ void __launch_bounds__(256) k_mc_centered(Geo g, const int* __restrict__ flags, CV3 vel, const float* __restrict__ orig,
                                                     const float* __restrict__ fwd, float* __restrict__ dst, float dt, float strength,
                                                     int clampMode, long long np) {
  IJK_FROM_BLOCK
  const bool fluid = fIsFluid(flags[idx]);
  const double s2 = (double)strength * 0.5;
  if (isBorder(g, i, j, k)) {
    // MacCormackCorrect is idx-mode over all cells; fwd and bwd are zero on the border (cleared temp grids)
#pragma unroll
    for (int c = 0; c < NC; c++) {
      float d = 0.0f;
      if (fluid) {
        const float o = orig[idx + c * np];
        if (NC == 1) d = (float)((double)0.0f + s2 * (double)(o - 0.0f));
        else         d = 0.0f + (float)(s2 * (double)(o - 0.0f));
      }
      dst[idx + c * np] = d;
    }
    return;
  }
  // backward step: SemiLagrange(flags, vel, bwd, fwd, -dt)
  float px, py, pz;
  tracePos<OT>(g, vel, i, j, k, idx, -dt, px, py, pz);
  float fw[NC], dv[NC];
#pragma unroll
  for (int c = 0; c < NC; c++) {
    fw[c] = fwd[idx + c * np];
    const float bw = sampleHi<OS, (NC == 3)>(fwd + c * np, g, px, py, pz);
    float d = fw[c];
    if (fluid) {
      const float o = orig[idx + c * np];
      if (NC == 1) d = (float)((double)d + s2 * (double)(o - bw));       // Real: whole update in double (advection.cpp:90)
      else         d = d + (float)(s2 * (double)(o - bw));               // Vec3: scalar*vector rounds to float first
    }
    dv[c] = d;
  }
  // MacCormackClamp (advection.cpp:242-268) with doClampComponent (:146-187)
  float cx, cy, cz;
  getCentered(vel, g, idx, cx, cy, cz);
  const float vx = cx * dt, vy = cy * dt, vz = cz * dt;
  const int ux = g.nx - 1, uy = g.ny - 1, uz = g.nz - 1;   // gridUpper
  float minv[NC], maxv[NC];
#pragma unroll
  for (int c = 0; c < NC; c++) { minv[c] = FLT_MAX; maxv[c] = -FLT_MAX; }
  bool haveFl = false;
  const int numPos = (clampMode == 1) ? 2 : 1;
  for (int l = 0; l < numPos; l++) {
    int cpx, cpy, cpz;
    if (l == 0) { cpx = (int)((float)i - vx); cpy = (int)((float)j - vy); cpz = (int)((float)k - vz); }
    else        { cpx = (int)((float)i + vx); cpy = (int)((float)j + vy); cpz = (int)((float)k + vz); }
    const int i0 = min(max(cpx, 0), ux - 1);
    const int j0 = min(max(cpy, 0), uy - 1);
    const int k0 = min(max(cpz, 0), g.is3D ? (uz - 1) : 1);   // 2-D upper bound is 1, as in the reference (:164)
    const long long b = (long long)i0 + (long long)g.nx * j0 + g.Z * k0;
    const int nk = g.is3D ? 2 : 1;
    for (int dk = 0; dk < nk; dk++)
      for (int dj = 0; dj < 2; dj++)
        for (int di = 0; di < 2; di++) {
          const long long n = b + di + (long long)g.nx * dj + g.Z * dk;
          if (CHECKFLAG(flags[n])) {
            haveFl = true;
#pragma unroll
            for (int c = 0; c < NC; c++) { const float o = orig[n + c * np]; if (o < minv[c]) minv[c] = o; if (o > maxv[c]) maxv[c] = o; }
          }
        }
  }
  if (!haveFl) {
#pragma unroll
    for (int c = 0; c < NC; c++) dv[c] = fw[c];
  } else if (clampMode == 1) {
#pragma unroll
    for (int c = 0; c < NC; c++) { if (dv[c] < minv[c]) dv[c] = minv[c]; else if (dv[c] > maxv[c]) dv[c] = maxv[c]; }
  } else {
    bool out = false;
#pragma unroll
    for (int c = 0; c < NC; c++) out |= (dv[c] < minv[c]) | (dv[c] > maxv[c]);
    if (out) {
#pragma unroll
      for (int c = 0; c < NC; c++) dv[c] = fw[c];
    }
  }
  if (clampMode == 1) {
    const int fx = (int)(((float)i + 0.5f) - vx), fy = (int)(((float)j + 0.5f) - vy), fz = (int)(((float)k + 0.5f) - vz);
    const int bx = (int)(((float)i + 0.5f) + vx), by = (int)(((float)j + 0.5f) + vy), bz = (int)(((float)k + 0.5f) + vz);
    bool revert = fx < 0 || fy < 0 || fz < 0 || bx < 0 || by < 0 || bz < 0 ||
                  fx > ux || fy > uy || ((fz > uz) && g.is3D) || bx > ux || by > uy || ((bz > uz) && g.is3D);
    if (!revert) {
      const int ff = flags[(long long)fx + (long long)g.nx * fy + g.Z * fz];
      const int fb = flags[(long long)bx + (long long)g.nx * by + g.Z * bz];
      revert = fIsObstacle(ff) || fIsObstacle(fb);
    }
    if (revert) {
#pragma unroll
      for (int c = 0; c < NC; c++) dv[c] = fw[c];
    }
  }
#pragma unroll
  for (int c = 0; c < NC; c++) dst[idx + c * np] = dv[c];
}

// ---- MAC grids ------------------------------------------------------------------------------
template <int C>
__device__ __forceinline__ void getAtMAC(const CV3& v, const Geo& g, long long idx, float& cx, float& cy, float& cz) {
  if (C == 0) getAtMACX(v, g, idx, cx, cy, cz);
  else if (C == 1) getAtMACY(v, g, idx, cx, cy, cz);
  else getAtMACZ(v, g, idx, cx, cy, cz);
}
__device__ __forceinline__ const float* compPtr(const CV3& v, int c) { return c == 0 ? v.x : (c == 1 ? v.y : v.z); }

// one component of SemiLagrangeMAC (advection.cpp:45-78)
template <int C, int OS, int OT>
__device__ __forceinline__ float slMacComp(const Geo& g, const CV3& vel, const CV3& src, int i, int j, int k, long long idx, float dt) {
  const float p0x = i + 0.5f, p0y = j + 0.5f, p0z = k + 0.5f;
  float px, py, pz;
  if (OT == 1) {
    float cx, cy, cz;
    getAtMAC<C>(vel, g, idx, cx, cy, cz);
    px = p0x - cx * dt; py = p0y - cy * dt; pz = p0z - cz * dt;
  } else {
    float cx, cy, cz;
    getAtMAC<C>(src, g, idx, cx, cy, cz);   // note: src, not vel (advection.cpp:62,66,70)
    const float q0x = (C == 0) ? (float)i : p0x, q0y = (C == 1) ? (float)j : p0y, q0z = (C == 2) ? (float)k : p0z;
    const float p1x = q0x - (cx * dt) * 0.5f, p1y = q0y - (cy * dt) * 0.5f, p1z = q0z - (cz * dt) * 0.5f;
    float vx, vy, vz;
    interpolMAC(src.x, src.y, src.z, g, p1x, p1y, p1z, vx, vy, vz);
    px = p0x - vx * dt; py = p0y - vy * dt; pz = p0z - vz * dt;
  }
  if (OS == 1) return interpol(compPtr(src, C), g, px, py, pz);
  // interpolCubicMAC(...)[C]: shift by +0.5 on the component's own axis (interpolHigh.h:174-181)
  if (C == 2 && g.Z == 0) return 0.0f;
  return interpolCubic<true>(compPtr(src, C), g, px + (C == 0 ? 0.5f : 0.0f), py + (C == 1 ? 0.5f : 0.0f), pz + (C == 2 ? 0.5f : 0.0f));
}

template <int OS, int OT>
__global__ void __launch_bounds__(256) k_sl_mac(Geo g, CV3 vel, CV3 src, V3 dst, float dt) {
  IJK_FROM_BLOCK
  if (isBorder(g, i, j, k)) { dst.x[idx] = 0.0f; dst.y[idx] = 0.0f; dst.z[idx] = 0.0f; return; }
  dst.x[idx] = slMacComp<0, OS, OT>(g, vel, src, i, j, k, idx, dt);
  dst.y[idx] = slMacComp<1, OS, OT>(g, vel, src, i, j, k, idx, dt);
  dst.z[idx] = slMacComp<2, OS, OT>(g, vel, src, i, j, k, idx, dt);
}

// doClampComponentMAC<c> (advection.cpp:192-236)
template <int C>
__device__ __forceinline__ float clampCompMAC(const Geo& g, const int* __restrict__ flags, float dst, const float* __restrict__ orig,
                                              float fwd, int i, int j, int k, long long idx, float vx, float vy, float vz, int clampMode) {
  if (clampMode == 2) {
    const long long nb = idx - (C == 0 ? 1 : (C == 1 ? (long long)g.nx : g.Z));
    if (!(CHECKFLAG(flags[idx]) && CHECKFLAG(flags[nb]))) return fwd;
  }
  float minv = FLT_MAX, maxv = -FLT_MAX;
  const int numPos = (clampMode == 1) ? 2 : 1;
  for (int l = 0; l < numPos; l++) {
    int cpx, cpy, cpz;
    if (l == 0) { cpx = (int)((float)i - vx); cpy = (int)((float)j - vy); cpz = (int)((float)k - vz); }
    else        { cpx = (int)((float)i + vx); cpy = (int)((float)j + vy); cpz = (int)((float)k + vz); }
    const int i0 = min(max(cpx, 0), g.nx - 2);
    const int j0 = min(max(cpy, 0), g.ny - 2);
    const int k0 = min(max(cpz, 0), g.is3D ? (g.nz - 2) : 0);
    const long long b = (long long)i0 + (long long)g.nx * j0 + g.Z * k0;
    const int nk = g.is3D ? 2 : 1;
    for (int dk = 0; dk < nk; dk++)
      for (int dj = 0; dj < 2; dj++)
        for (int di = 0; di < 2; di++) {
          const float o = orig[b + di + (long long)g.nx * dj + g.Z * dk];
          if (o < minv) minv = o;
          if (o > maxv) maxv = o;
        }
  }
  if (clampMode == 1) { if (dst < minv) return minv; if (dst > maxv) return maxv; return dst; }
  if (dst < minv || dst > maxv) return fwd;
  return dst;
}

template <int C, int OS, int OT>
__device__ __forceinline__ float mcMacComp(const Geo& g, const int* __restrict__ flags, const CV3& vel, const CV3& orig, const CV3& fwd,
                                           int i, int j, int k, long long idx, float dt, double s2, int clampMode, bool skip, bool border) {
  const float* fwdc = compPtr(fwd, C);
  const float* origc = compPtr(orig, C);
  if (border) {
    // fwd == bwd == 0 on the border; MacCormackCorrectMAC still runs there (KERNEL() without bnd)
    return skip ? 0.0f : (float)((double)0.0f + s2 * (double)(origc[idx] - 0.0f));
  }
  const float fw = fwdc[idx];
  float d = fw;
  if (!skip) {
    const float bw = slMacComp<C, OS, OT>(g, vel, fwd, i, j, k, idx, -dt);
    d = (float)((double)fw + s2 * (double)(origc[idx] - bw));    // advection.cpp:113
  }
  if (C == 2 && !g.is3D) return d;                             // z component is not clamped in 2-D (:281-282)
  float cx, cy, cz;
  getAtMAC<C>(vel, g, idx, cx, cy, cz);
  return clampCompMAC<C>(g, flags, d, origc, fw, i, j, k, idx, cx * dt, cy * dt, cz * dt, clampMode);
}

template <int OS, int OT>
__global__ void __launch_bounds__(256) k_mc_mac(Geo g, const int* __restrict__ flags, CV3 vel, CV3 orig, CV3 fwd, V3 dst, float dt,
                                                float strength, int clampMode) {
  IJK_FROM_BLOCK
  const int f = flags[idx];
  const bool nf = !fIsFluid(f);
  // skip[] of MacCormackCorrectMAC (advection.cpp:99-106), isMAC == true
  bool sx = nf, sy = nf, sz = nf;
  if (i > 0 && !fIsFluid(flags[idx - 1])) sx = true;
  if (j > 0 && !fIsFluid(flags[idx - g.nx])) sy = true;
  if (k > 0 && !fIsFluid(flags[idx - g.Z])) sz = true;
  const bool border = isBorder(g, i, j, k);
  const double s2 = (double)strength * 0.5;
  dst.x[idx] = mcMacComp<0, OS, OT>(g, flags, vel, orig, fwd, i, j, k, idx, dt, s2, clampMode, sx, border);
  dst.y[idx] = mcMacComp<1, OS, OT>(g, flags, vel, orig, fwd, i, j, k, idx, dt, s2, clampMode, sy, border);
  dst.z[idx] = mcMacComp<2, OS, OT>(g, flags, vel, orig, fwd, i, j, k, idx, dt, s2, clampMode, sz, border);
}

__global__ void k_scan_outflow(const int* __restrict__ flags, long long n, int* anyOutflow) {   // flags: the rank's slab
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  bool any = false;
  for (; i < n; i += st) any |= fIsOutflow(flags[i]);
  if (any) *anyOutflow = 1;
}

// ---- outflow BC (advection.cpp:327-392) --------------------------------------------------------
__device__ __forceinline__ bool inBounds(const Geo& g, int x, int y, int z) {
  return x >= 0 && y >= 0 && z >= 0 && x < g.nx && y < g.ny && z < g.nz;
}

__global__ void k_outflow_extrapolate(Geo g, const int* __restrict__ flags, CV3 vel, V3 velDst, CV3 velPrev, float timeStep) {
  IJK_FROM_BLOCK
  if (!fIsOutflow(flags[idx])) return;
  // getBulkVel (:327-344)
  float ax = 0.f, ay = 0.f, az = 0.f;
  int count = 0;
  const int nmax = g.is3D ? 1 : 0;
  for (int n = -nmax; n <= nmax; n++)
    for (int m = -1; m <= 1; m++)
      for (int l = -1; l <= 1; l++) {
        if (inBounds(g, i + l, j + m, k + n)) {
          const long long q = idx + l + (long long)g.nx * m + g.Z * n;
          const int fq = flags[q];
          if (fIsFluid(fq) || fIsOutflow(fq)) { ax += vel.x[q]; ay += vel.y[q]; az += vel.z[q]; count++; }
        }
      }
  if (count > 0) { ax = ax / (float)count; ay = ay / (float)count; az = az / (float)count; }
  const float bulk[3] = { ax, ay, az };
  const int dim = g.is3D ? 3 : 2;
  const int cur[3] = { i, j, k };
  const long long strides[3] = { 1, (long long)g.nx, g.Z };
  const float v0x = vel.x[idx], v0y = vel.y[idx], v0z = vel.z[idx];
  const float dvx = v0x - velPrev.x[idx], dvy = v0y - velPrev.y[idx], dvz = v0z - velPrev.z[idx];
  float rx = 0.f, ry = 0.f, rz = 0.f;
  int cnt = 0;
  for (int c = 0; c < dim; c++) {
    const float factor = timeStep * fmaxf(1.0f, bulk[c]);
    int flLow[3] = { cur[0], cur[1], cur[2] }, flUp[3] = { cur[0], cur[1], cur[2] };
    flLow[c] = cur[c] - 1; flUp[c] = cur[c] + 1;
    const long long low = idx - strides[c], up = idx + strides[c];   // 'low'/'up' stay one cell away (:358-359)
    for (int d = 0; d < 2; d++) {
      const bool fromLower = inBounds(g, flLow[0], flLow[1], flLow[2]) &&
                             fIsFluid(flags[(long long)flLow[0] + (long long)g.nx * flLow[1] + g.Z * flLow[2]]);
      const bool fromUpper = inBounds(g, flUp[0], flUp[1], flUp[2]) &&
                             fIsFluid(flags[(long long)flUp[0] + (long long)g.nx * flUp[1] + g.Z * flUp[2]]);
      if (fromLower || fromUpper) {
        if (fromLower) { rx += (dvx / factor) + vel.x[low]; ry += (dvy / factor) + vel.y[low]; rz += (dvz / factor) + vel.z[low]; cnt++; }
        if (fromUpper) { rx += (dvx / factor) + vel.x[up];  ry += (dvy / factor) + vel.y[up];  rz += (dvz / factor) + vel.z[up];  cnt++; }
        break;
      }
      flLow[c]--; flUp[c]++;
    }
  }
  if (cnt > 0) { rx = rx / (float)cnt; ry = ry / (float)cnt; rz = rz / (float)cnt; }
  velDst.x[idx] = rx; velDst.y[idx] = ry; velDst.z[idx] = rz;
}

__global__ void k_outflow_copy(Geo g, const int* __restrict__ flags, CV3 velDst, V3 vel) {
  IJK_FROM_BLOCK
  if (fIsOutflow(flags[idx])) { vel.x[idx] = velDst.x[idx]; vel.y[idx] = velDst.y[idx]; vel.z[idx] = velDst.z[idx]; }
}


// ---- host side ------------------------------------------------------------------------------------
bool flagsHaveOutflow(const Grid* flags) {
  if (flags->outflowState >= 0) return flags->outflowState != 0;
  Solver* s = flags->s;
  s->clearShared(s->devFlagsScratch, sizeof(int));
  MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH(s, k_scan_outflow, s->numSMs * 8, 256, flags->di() + sp.off, sp.n, s->devFlagsScratch); }
  s->rankSync();
  s->readSmall(s->hostInts, s->devFlagsScratch, (int)sizeof(int));
  flags->outflowState = s->hostInts[0] ? 1 : 0;
  return flags->outflowState != 0;
}

// applyOutflowBC(flags, vel, velPrev, dt)  (advection.cpp:388-392)
static void applyOutflowBC(Solver* s, const Grid* flags, V3 vel, CV3 velPrev, float dt) {
  if (!flagsHaveOutflow(flags)) return;
  TmpVec velDst(s);
  const float timeStep = (float)fmax(1.0, (double)dt * 4);   // max(1.0, timeStep*4) evaluated in double, passed as Real
  CV3 cvel; cvel.x = vel.x; cvel.y = vel.y; cvel.z = vel.z;
  MF_LAUNCH_IJK(s, k_outflow_extrapolate, flags->di(), cvel, v3(velDst), velPrev, timeStep);
  MF_LAUNCH_IJK(s, k_outflow_copy, flags->di(), cv3(velDst), vel);
}

// the 32-bit-index kernels of advect_fast.cuh: 3-D, linear interpolation, first-order trace (MF_ADVECT_GENERIC=1: tests compare
// them with the generic kernels)
static bool fastPath(const Solver* s, int OS, int OT) {
  return s->is3D && OS == 1 && OT == 1 && s->N < (1LL << 31) && !getenv("MF_ADVECT_GENERIC");
}

template <int NC, int OS, int OT>
static void advectCentered(Solver* s, const Grid* flags, const Grid* vel, Grid* grid, int order, float strength, int clampMode) {
  const long long np = s->Np;
  const float dt = s->dt;
  const bool fast = NC == 1 && fastPath(s, OS, OT);
  float* fwd = (NC == 1) ? s->getReal() : s->getVec();
  if (fast) MF_LAUNCH_IJK_PDL(s, k_sl_f, cv3(vel), grid->d, fwd, dt);
  else MF_LAUNCH_IJK(s, (k_sl<NC, OS, OT>), cv3(vel), grid->d, fwd, dt, np);
  if (order == 1) {
    float* old = grid->d; grid->d = fwd; fwd = old;       // orig.swap(fwd)
  } else {
    float* nw = (NC == 1) ? s->getReal() : s->getVec();
    if (fast && clampMode == 1) MF_LAUNCH_IJK_PDL(s, k_mc_centered_f<1>, flags->di(), cv3(vel), grid->d, fwd, nw, dt, strength, mcPow2Scale(strength));
    else if (fast) MF_LAUNCH_IJK_PDL(s, k_mc_centered_f<2>, flags->di(), cv3(vel), grid->d, fwd, nw, dt, strength, mcPow2Scale(strength));
    else MF_LAUNCH_IJK(s, (k_mc_centered<NC, OS, OT>), flags->di(), cv3(vel), grid->d, fwd, nw, dt, strength, clampMode, np);
    float* old = grid->d; grid->d = nw;                   // orig.swap(newGrid)
    if (NC == 1) s->putReal(old); else s->putVec(old);
  }
  if (NC == 1) s->putReal(fwd); else s->putVec(fwd);
}

static bool tmaEnabled() { const char* e = getenv("MF_ADVECT_TMA"); return e && atoi(e) != 0; }
static void launchSlMacTma(Solver* s, const Grid* vel, Grid* fw, float dt) {
  static bool attr = false;
  if (!attr) { CUDA_CHECK(cudaFuncSetAttribute(k_sl_mac_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tmaSmemBytes())); attr = true; }
  TmaMaps3 maps;
  for (int c = 0; c < 3; c++) maps.m[c] = tmaPlaneMap(s, vel->comp(c));
  const int zc = getenv("MF_TMA_ZC") ? atoi(getenv("MF_TMA_ZC")) : 32;
  const Geo g = geo(s);
  const int tiles = ((s->nx + TMA_TX - 1) / TMA_TX) * ((s->ny + TMA_TY - 1) / TMA_TY), chunks = (s->nz + zc - 1) / zc;
  ProfScope ps(s, "k_sl_mac_tma");
  k_sl_mac_tma<<<tiles * chunks, 256, tmaSmemBytes(), s->stream>>>(g, maps, cv3(vel), v3(fw), dt, zc);
  s->launches++;
}

template <int OS, int OT>
static void advectMAC(Solver* s, const Grid* flags, const Grid* vel, Grid* grid, int order, float strength, int clampMode) {
  const float dt = s->dt;
  const bool self = (vel == grid) || (vel->d == grid->d);
  float* fwd = s->getVec();
  Grid fw = *grid; fw.d = fwd;
  const bool fast = fastPath(s, OS, OT);
  // MF_ADVECT_TMA=1: the TMA-staged variant (advect_tma.cuh) for self-advection on a single GPU
  if (fast && self && tmaUsable(s) && tmaEnabled()) launchSlMacTma(s, vel, &fw, dt);
  else if (fast) MF_LAUNCH_IJK_PDL(s, k_sl_mac_f, cv3(vel), cv3(grid), v3(&fw), dt);
  else MF_LAUNCH_IJK(s, (k_sl_mac<OS, OT>), cv3(vel), cv3(grid), v3(&fw), dt);
  if (order == 1) {
    applyOutflowBC(s, flags, v3(&fw), cv3(grid), dt);
    float* old = grid->d; grid->d = fwd; fwd = old;
  } else {
    float* nw = s->getVec();
    Grid ng = *grid; ng.d = nw;
    if (fast && clampMode == 1) MF_LAUNCH_IJK_PDL(s, k_mc_mac_f<1>, flags->di(), cv3(vel), cv3(grid), cv3(&fw), v3(&ng), dt, strength, mcPow2Scale(strength));
    else if (fast) MF_LAUNCH_IJK_PDL(s, k_mc_mac_f<2>, flags->di(), cv3(vel), cv3(grid), cv3(&fw), v3(&ng), dt, strength, mcPow2Scale(strength));
    else MF_LAUNCH_IJK(s, (k_mc_mac<OS, OT>), flags->di(), cv3(vel), cv3(grid), cv3(&fw), v3(&ng), dt, strength, clampMode);
    applyOutflowBC(s, flags, v3(&ng), cv3(grid), dt);
    float* old = grid->d; grid->d = nw;
    s->putVec(old);
  }
  s->putVec(fwd);
  (void)self;
}

} // namespace mf

using namespace mf;

extern "C" int mf_advect_semi_lagrange(const mf_grid* flagsp, const mf_grid* velp, mf_grid* gridp,
                                       int order, float strength, int orderSpace, int clampMode, int orderTrace) {
  MF_API_BEGIN
  REQ_GRID(flagsp, G(flagsp)->isInt(), "flags", "advectSemiLagrange");
  REQ_GRID(velp, G(velp)->isMAC(), "vel", "advectSemiLagrange");
  mfAssert(gridp, "advectSemiLagrange: argument 'grid' is missing");
  const Grid* flags = G(flagsp); const Grid* vel = G(velp); Grid* grid = G(gridp);
  Solver* s = flags->s; s->use();
  check_same_solver(flags, vel, "advectSemiLagrange"); check_same_solver(flags, grid, "advectSemiLagrange");
  mfAssert(order == 1 || order == 2, "AdvectSemiLagrange: Only order 1 (regular SL) and 2 (MacCormack) supported");
  mfAssert(orderSpace == 1 || orderSpace == 2, "Unknown interpolation order " << orderSpace);
  mfAssert(orderTrace == 1 || orderTrace == 2, "Unknown backtracing order " << orderTrace);
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "advectSemiLagrange: grid too small");
#define DISPATCH_OS_OT(FN, ...) \
  do { if (orderSpace == 1) { if (orderTrace == 1) FN<__VA_ARGS__ 1, 1>(s, flags, vel, grid, order, strength, clampMode); \
                              else FN<__VA_ARGS__ 1, 2>(s, flags, vel, grid, order, strength, clampMode); } \
       else { if (orderTrace == 1) FN<__VA_ARGS__ 2, 1>(s, flags, vel, grid, order, strength, clampMode); \
              else FN<__VA_ARGS__ 2, 2>(s, flags, vel, grid, order, strength, clampMode); } } while (0)
  // same dispatch order as advection.cpp:450-460: Real, then MAC, then Vec3
  if (grid->isReal()) { DISPATCH_OS_OT(advectCentered, 1, ); }
  else if (grid->isMAC()) { DISPATCH_OS_OT(advectMAC, ); }
  else if (grid->isVec()) { DISPATCH_OS_OT(advectCentered, 3, ); }
  else mfErr("AdvectSemiLagrange: Grid Type is not supported (only Real, Vec3, MAC, Levelset)");
  MF_LAUNCH_CHECK();
  MF_API_END
}
