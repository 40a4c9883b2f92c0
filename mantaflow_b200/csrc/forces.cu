// forces.cu -- FlagGrid setup, wall boundary conditions and body forces.
// Reference: source/grid.cpp (initDomain :798, initBoundaries :890, fillGrid :922),
//            source/plugin/extforces.cpp (KnApplyForce :46, KnAddBuoyancy :73, setOpenBound :106,
//            resetOutflow :134, KnSetInflow :164, KnSetWallBcs :187, KnSetWallBcsFrac :240),
//            source/shapes.cpp (ApplyShapeToGrid :40, Box::isInside :151, Cylinder::isInside :324).
#include "common.cuh"
#include "sample.cuh"
#include "quad.cuh"
#include <cstring>
#include <cmath>

namespace mf {

#define IJK_FROM_BLOCK \
  const int i = blockIdx.x * blockDim.x + threadIdx.x; \
  const int j = blockIdx.y * blockDim.y + threadIdx.y; \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0; \
  if (i >= g.nx || j >= g.ny || k >= g.k1) return; \
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;


__device__ __forceinline__ bool interior1(const Geo& g, int i, int j, int k) {
  return i >= 1 && j >= 1 && i < g.nx - 1 && j < g.ny - 1 && (!g.is3D || (k >= 1 && k < g.nz - 1));
}

// ---- flags -------------------------------------------------------------------------------------
struct Types6 { int t[6]; };

// setConst(TypeEmpty) + initBoundaries (grid.cpp:886-908): later sides overwrite earlier ones
__global__ void k_init_domain(Geo g, int* __restrict__ flags, int w, Types6 types) {
  IJK_FROM_BLOCK
  int f = MF_EMPTY;
  if (i <= w) f = types.t[0];
  if (i >= g.nx - 1 - w) f = types.t[1];
  if (j <= w) f = types.t[2];
  if (j >= g.ny - 1 - w) f = types.t[3];
  if (g.is3D) {
    if (k <= w) f = types.t[4];
    if (k >= g.nz - 1 - w) f = types.t[5];
  }
  flags[idx] = f;
}
// phiWalls (grid.cpp:756-796): phi = min(dist to each wall - 0.5 - w ...), computed in double like the reference
__global__ void k_init_phi_walls(Geo g, float* __restrict__ phi, int w, int wallMask) {
  IJK_FROM_BLOCK
  float p = 1000000000.0f;
  if (wallMask & 1)  p = (float)fmin((double)i - 0.5 - (double)w, (double)p);
  if (wallMask & 2)  p = (float)fmin((double)g.nx - (double)i - 1.5 - (double)w, (double)p);
  if (wallMask & 4)  p = (float)fmin((double)j - 0.5 - (double)w, (double)p);
  if (wallMask & 8)  p = (float)fmin((double)g.ny - (double)j - 1.5 - (double)w, (double)p);
  if (wallMask & 16) p = (float)fmin((double)k - 0.5 - (double)w, (double)p);
  if (wallMask & 32) p = (float)fmin((double)g.nz - (double)k - 1.5 - (double)w, (double)p);
  phi[idx] = p;
}
__global__ void k_fill_grid(int* __restrict__ flags, int type, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    const int f = flags[i];
    if ((f & MF_OBSTACLE) == 0 && (f & MF_INFLOW) == 0 && (f & MF_OUTFLOW) == 0 && (f & MF_OPEN) == 0)
      flags[i] = (f & ~(MF_EMPTY | MF_FLUID)) | type;
  }
}
__global__ void k_count_cells(Geo g, const int* __restrict__ flags, int flag, int bnd, int* out) {
  IJK_FROM_BLOCK
  const bool in = i >= bnd && j >= bnd && i < g.nx - bnd && j < g.ny - bnd && (!g.is3D || (k >= bnd && k < g.nz - bnd));
  const bool hit = in && ((flags[idx] & flag) != 0);
  const unsigned int m = __ballot_sync(__activemask(), hit);
  if (hit && (threadIdx.x & 31) == (__ffs(m) - 1)) atomicAdd(out, __popc(m));
}
// setOpenBound (extforces.cpp:106-131)
__global__ void k_set_open_bound(Geo g, int* __restrict__ flags, int bWidth, int lo, int up, int type) {
  IJK_FROM_BLOCK
  const bool loX = (lo & 1) && i <= bWidth, loY = (lo & 2) && j <= bWidth;
  const bool upX = (up & 1) && i >= g.nx - bWidth - 1, upY = (up & 2) && j >= g.ny - bWidth - 1;
  const bool innerI = i > bWidth && i < g.nx - bWidth - 1, innerJ = j > bWidth && j < g.ny - bWidth - 1;
  const int f = flags[idx];
  if (!g.is3D && (loX || upX || loY || upY)) {
    if ((loX || upX || innerI) && (loY || upY || innerJ) && fIsObstacle(f)) flags[idx] = type;
  } else {
    const bool loZ = (lo & 4) && k <= bWidth, upZ = (up & 4) && k >= g.nz - bWidth - 1;
    const bool innerK = k > bWidth && k < g.nz - bWidth - 1;
    if (loX || upX || loY || upY || loZ || upZ) {
      if ((loX || upX || innerI) && (loY || upY || innerJ) && (loZ || upZ || innerK) && fIsObstacle(f)) flags[idx] = type;
    }
  }
}
__global__ void k_reset_outflow(int* __restrict__ flags, float* phi, float* real, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    const int f = flags[i];
    if (fIsOutflow(f)) {
      flags[i] = (f | MF_EMPTY) & ~MF_FLUID;
      if (phi) phi[i] = 0.5f;
      if (real) real[i] = 0.0f;
    }
  }
}

// ---- shapes ---------------------------------------------------------------------------------------
template <class T>
__global__ void k_apply_box(Geo g, T* __restrict__ d, float x0, float y0, float z0, float x1, float y1, float z1, T value) {
  IJK_FROM_BLOCK
  const float px = i + 0.5f, py = j + 0.5f, pz = k + 0.5f;
  if (px >= x0 && py >= y0 && pz >= z0 && px <= x1 && py <= y1 && pz <= z1) d[idx] = value;
}
__global__ void k_apply_cylinder(Geo g, float* __restrict__ d, float cx, float cy, float cz, float radius, float zx, float zy, float zz,
                                 float zlen, float value) {
  IJK_FROM_BLOCK
  const float dx = (i + 0.5f) - cx, dy = (j + 0.5f) - cy, dz = (k + 0.5f) - cz;
  const float z = dx * zx + dy * zy + dz * zz;
  if (fabsf(z) > zlen) return;
  const float r2 = (dx * dx + dy * dy + dz * dz) - z * z;
  if (r2 < radius * radius) d[idx] = value;
}

// ---- forces ---------------------------------------------------------------------------------------
// vel + (0.5 * strength) * factorSum as the reference evaluates it: a double product and sum, rounded once.  For a zero strength
// component the float expression gives the same bits (the product is a zero of the same sign or NaN, the sum is exact), and
// two of three components of a gravity vector are zero: they stay off the FP64 pipe and its conversions.
__device__ __forceinline__ float buoy(float v, float strength, float factorSum) {
  if (strength == 0.f) return v + (0.5f * strength) * factorSum;
  return (float)((double)v + (0.5 * (double)strength) * (double)factorSum);
}
// KnAddBuoyancy (extforces.cpp:73-81): (0.5*strength) is a double product, the update is done in double.
// W x-consecutive cells per thread (quad.cuh); a component's cells are only stored when one of them changed.
template <int W>
__global__ void __launch_bounds__(256) k_add_buoyancy(Geo g, const int* __restrict__ flags, const float* __restrict__ factor, V3 vel,
                                                      float sx, float sy, float sz) {
  MF_PDL_PROLOGUE
  MF_CELLS_W(g, W)
  if (j < 1 || j >= g.ny - 1 || (g.is3D && (k < 1 || k >= g.nz - 1))) return;     // bnd = 1 rows only (warp-uniform)
  const long long Y = g.nx, Z = g.Z;
  // all loads first (flags, factor, velocity: one round trip to memory), the shuffles of the x - 1 neighbours behind them
  // (quad.cuh); rows j = 0 / k = 0 never reach this point, so idx - Y / idx - Z exist
  int f[W], fx[W], fy[W], fz[W];
  float d0[W], dx[W], dy[W], dz[W];
  float vx[W], vy[W], vz[W];
  ldW<W>(f, flags, idx); ldW<W>(fy, flags, idx - Y);
  if (g.is3D) ldW<W>(fz, flags, idx - Z);
  ldW<W>(d0, factor, idx); ldW<W>(dy, factor, idx - Y);
  if (g.is3D) ldW<W>(dz, factor, idx - Z);
  ldW<W>(vx, vel.x, idx); ldW<W>(vy, vel.y, idx);
  if (g.is3D) ldW<W>(vz, vel.z, idx);
  if constexpr (W > 1) {
    const int fe = xEdgeMinus<W, int>(flags, idx, i, 0);
    const float de = xEdgeMinus<W, float>(factor, idx, i, 0.f);
    xMinusWS<W, int>(fx, f, fe); xMinusWS<W, float>(dx, d0, de);
  } else { xMinusW<W, int>(fx, f, flags, idx, i, 0); xMinusW<W, float>(dx, d0, factor, idx, i, 0.f); }
  if (!g.is3D) { for (int c = 0; c < W; c++) { fz[c] = 0; dz[c] = 0.f; vz[c] = 0.f; } }
  if (!on) return;
  bool chx = false, chy = false, chz = false;
#pragma unroll
  for (int c = 0; c < W; c++) {
    const bool cell = (i + c >= 1) && (i + c < g.nx - 1) && fIsFluid(f[c]);
    if (cell && fIsFluid(fx[c])) {
      const float o = vx[c];
      vx[c] = buoy(o, sx, d0[c] + dx[c]);
      chx |= __float_as_uint(vx[c]) != __float_as_uint(o);
    }
    if (cell && fIsFluid(fy[c])) {
      const float o = vy[c];
      vy[c] = buoy(o, sy, d0[c] + dy[c]);
      chy |= __float_as_uint(vy[c]) != __float_as_uint(o);
    }
    if (cell && g.is3D && fIsFluid(fz[c])) {
      const float o = vz[c];
      vz[c] = buoy(o, sz, d0[c] + dz[c]);
      chz |= __float_as_uint(vz[c]) != __float_as_uint(o);
    }
  }
  // a component is stored only where its bits changed: a zero strength component leaves them as they are (except -0 -> +0
  // and non-finite factors)
  if (chx) stW<W>(vel.x, idx, vx);
  if (chy) stW<W>(vel.y, idx, vy);
  if (chz) stW<W>(vel.z, idx, vz);
}
// KnApplyForce (extforces.cpp:46-59), additive flag as in addGravity
__global__ void __launch_bounds__(256) k_apply_force(Geo g, const int* __restrict__ flags, V3 vel, float fx, float fy, float fz,
                                                     const float* __restrict__ exclude, int additive) {
  IJK_FROM_BLOCK
  if (!interior1(g, i, j, k)) return;
  const int f = flags[idx];
  const bool curFluid = fIsFluid(f), curEmpty = fIsEmpty(f);
  if (!curFluid && !curEmpty) return;
  if (exclude && exclude[idx] < 0.0f) return;
  const int fxm = flags[idx - 1], fym = flags[idx - g.nx];
  if (fIsFluid(fxm) || (curFluid && fIsEmpty(fxm))) vel.x[idx] = additive ? vel.x[idx] + fx : fx;
  if (fIsFluid(fym) || (curFluid && fIsEmpty(fym))) vel.y[idx] = additive ? vel.y[idx] + fy : fy;
  if (g.is3D) {
    const int fzm = flags[idx - g.Z];
    if (fIsFluid(fzm) || (curFluid && fIsEmpty(fzm))) vel.z[idx] = additive ? vel.z[idx] + fz : fz;
  }
}
// KnSetInflow (extforces.cpp:164-168)
__global__ void k_set_inflow(Geo g, V3 vel, int dim, int p0, float vx, float vy, float vz) {
  IJK_FROM_BLOCK
  const int p = dim == 0 ? i : (dim == 1 ? j : k);
  if (p == p0 || p == p0 + 1) { vel.x[idx] = vx; vel.y[idx] = vy; vel.z[idx] = vz; }
}

// KnSetWallBcs (extforces.cpp:187-237): all cells, in place.  W x-consecutive cells per thread (quad.cuh); the kernel streams
// the flags and touches a velocity component only where a wall condition applies (a thin shell of the domain).
// wallBcsApply: the rules for one group of W cells whose flags and the flags of the six neighbours are in registers.
template <int W>
__device__ __forceinline__ void wallBcsApply(const Geo& g, int i, int j, int k, long long idx, const int (&f)[W], const int (&fxm)[W], const int (&fxp)[W],
                                             const int (&fym)[W], const int (&fyp)[W], const int (&fzm)[W], const int (&fzp)[W],
                                             V3 vel, CV3 obvel, int haveObvel) {
  // what happens to each component of each cell: 0 keep, 1 set to the obstacle velocity, 2 set to zero (applied in this order)
  int ax[W], ay[W], az[W];
  bool anyx = false, anyy = false, anyz = false;
#pragma unroll
  for (int c = 0; c < W; c++) {
    ax[c] = ay[c] = az[c] = 0;
    if (g.is3D && !((f[c] | fxm[c] | fxp[c] | fym[c] | fyp[c] | fzm[c] | fzp[c]) & (MF_OBSTACLE | MF_STICK))) continue;
    const bool curFluid = fIsFluid(f[c]), curObs = fIsObstacle(f[c]);
    if (!curFluid && !curObs) continue;
    // (neighbours outside the grid were loaded as 0: no obstacle, no fluid, no stick bit -- the same as the index guards)
    if (i + c > 0 && (fIsObstacle(fxm[c]) || (curObs && fIsFluid(fxm[c])))) ax[c] = 1;
    if (j > 0 && (fIsObstacle(fym[c]) || (curObs && fIsFluid(fym[c])))) ay[c] = 1;
    if (!g.is3D) az[c] = 2;
    else if (k > 0 && (fIsObstacle(fzm[c]) || (curObs && fIsFluid(fzm[c])))) az[c] = 1;
    if (curFluid) {
      if (fIsStick(fxm[c]) || fIsStick(fxp[c])) { ay[c] = 2; az[c] = 2; }
      if (fIsStick(fym[c]) || fIsStick(fyp[c])) { ax[c] = 2; az[c] = 2; }
      if (g.is3D && (fIsStick(fzm[c]) || fIsStick(fzp[c]))) { ax[c] = 2; ay[c] = 2; }
    }
    anyx |= ax[c] != 0; anyy |= ay[c] != 0; anyz |= az[c] != 0;
  }
  float v[W], o[W];
  if (anyx) {
    ldW<W>(v, vel.x, idx);
    if (haveObvel) ldW<W>(o, obvel.x, idx); else { for (int c = 0; c < W; c++) o[c] = 0.f; }
#pragma unroll
    for (int c = 0; c < W; c++) v[c] = ax[c] == 0 ? v[c] : (ax[c] == 1 ? o[c] : 0.f);
    stW<W>(vel.x, idx, v);
  }
  if (anyy) {
    ldW<W>(v, vel.y, idx);
    if (haveObvel) ldW<W>(o, obvel.y, idx); else { for (int c = 0; c < W; c++) o[c] = 0.f; }
#pragma unroll
    for (int c = 0; c < W; c++) v[c] = ay[c] == 0 ? v[c] : (ay[c] == 1 ? o[c] : 0.f);
    stW<W>(vel.y, idx, v);
  }
  if (anyz) {
    ldW<W>(v, vel.z, idx);
    if (haveObvel && g.is3D) ldW<W>(o, obvel.z, idx); else { for (int c = 0; c < W; c++) o[c] = 0.f; }
#pragma unroll
    for (int c = 0; c < W; c++) v[c] = az[c] == 0 ? v[c] : (az[c] == 1 ? o[c] : 0.f);
    stW<W>(vel.z, idx, v);
  }
}
// split != 0 (3-D, W = 4): the groups at the two x walls are left to k_set_wall_bcs_edge.  A row of 4-cell groups is a whole
// number of warps whose first / last lane sits at a wall, so with those groups in the same launch NO warp could take the early
// exit below and the kernel was bound by the rules' instructions (380 per warp) instead of the flag stream.
template <int W>
__global__ void __launch_bounds__(256) k_set_wall_bcs(Geo g, const int* __restrict__ flags, V3 vel, CV3 obvel, int haveObvel, int split) {
  MF_PDL_PROLOGUE
  MF_CELLS_W(g, W)
  const long long Y = g.nx, Z = g.Z;
  int f[W], fxm[W], fxp[W], fym[W], fyp[W], fzm[W], fzp[W];
  // all loads first, the shuffles of the x neighbours behind them (quad.cuh)
  ldW<W>(f, flags, idx);
  if (j > 0) ldW<W>(fym, flags, idx - Y); else { for (int c = 0; c < W; c++) fym[c] = 0; }
  if (j < g.ny - 1) ldW<W>(fyp, flags, idx + Y); else { for (int c = 0; c < W; c++) fyp[c] = 0; }
  if (g.is3D && k > 0) ldW<W>(fzm, flags, idx - Z); else { for (int c = 0; c < W; c++) fzm[c] = 0; }
  if (g.is3D && k < g.nz - 1) ldW<W>(fzp, flags, idx + Z); else { for (int c = 0; c < W; c++) fzp[c] = 0; }
  if constexpr (W > 1) {
    const int le = xEdgeMinus<W, int>(flags, idx, i, 0), re = xEdgePlus<W, int>(flags, idx, i, g.nx, 0);
    xMinusWS<W, int>(fxm, f, le); xPlusWS<W, int>(fxp, f, re, i, g.nx);
  } else { xMinusW<W, int>(fxm, f, flags, idx, i, 0); xPlusW<W, int>(fxp, f, flags, idx, i, g.nx, 0); }
  if (!on) return;
  if (split && (i == 0 || i + W >= g.nx)) return;                      // k_set_wall_bcs_edge
  if (g.is3D) {
    // every rule needs an obstacle or a stick bit in the cell or one of its six neighbours: the bulk of the domain leaves here
    int any = 0;
#pragma unroll
    for (int c = 0; c < W; c++) any |= f[c] | fxm[c] | fxp[c] | fym[c] | fyp[c] | fzm[c] | fzp[c];
    if (!(any & (MF_OBSTACLE | MF_STICK))) return;
  }
  wallBcsApply<W>(g, i, j, k, idx, f, fxm, fxp, fym, fyp, fzm, fzp, vel, obvel, haveObvel);
}
// the first and the last group of every row (blockDim.x == 2: side), see above
template <int W>
__global__ void __launch_bounds__(256) k_set_wall_bcs_edge(Geo g, const int* __restrict__ flags, V3 vel, CV3 obvel, int haveObvel) {
  const int side = threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  if (j >= g.ny || k >= g.k1) return;
  if (side == 1 && g.nx <= W) return;                                   // one group per row: side 0 has it
  const int i = side == 0 ? 0 : g.nx - W;
  const long long Y = g.nx, Z = g.Z;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  int f[W], fxm[W], fxp[W], fym[W], fyp[W], fzm[W], fzp[W];
  ldW<W>(f, flags, idx);
  if (j > 0) ldW<W>(fym, flags, idx - Y); else { for (int c = 0; c < W; c++) fym[c] = 0; }
  if (j < g.ny - 1) ldW<W>(fyp, flags, idx + Y); else { for (int c = 0; c < W; c++) fyp[c] = 0; }
  if (g.is3D && k > 0) ldW<W>(fzm, flags, idx - Z); else { for (int c = 0; c < W; c++) fzm[c] = 0; }
  if (g.is3D && k < g.nz - 1) ldW<W>(fzp, flags, idx + Z); else { for (int c = 0; c < W; c++) fzp[c] = 0; }
  const int le = (i > 0) ? flags[idx - 1] : 0, re = (i + W < g.nx) ? flags[idx + W] : 0;
#pragma unroll
  for (int c = 0; c < W; c++) { fxm[c] = c > 0 ? f[c - 1] : le; fxp[c] = c + 1 < W ? f[c + 1] : re; }
  wallBcsApply<W>(g, i, j, k, idx, f, fxm, fxp, fym, fyp, fzm, fzp, vel, obvel, haveObvel);
}

// KnSetWallBcsFrac (extforces.cpp:240-323): writes velTarget for every cell
__global__ void __launch_bounds__(256) k_set_wall_bcs_frac(Geo g, const int* __restrict__ flags, CV3 vel, V3 tgt, const float* __restrict__ phiObs) {
  IJK_FROM_BLOCK
  const int f = flags[idx];
  const bool curFluid = fIsFluid(f), curObs = fIsObstacle(f);
  float tx = vel.x[idx], ty = vel.y[idx], tz = vel.z[idx];
  if ((curFluid || curObs) && interior1(g, i, j, k)) {
    const long long X = 1, Y = g.nx, Z = g.Z;
    const float* p = phiObs + idx;
    if (curObs | fIsObstacle(flags[idx - X])) {
      float dx = 0.f, dy = 0.f, dz = 0.f;
      const float tmp1 = (p[0] + p[-X]) * .5f;
      float tmp2 = (p[Y] + p[-X + Y]) * .5f;
      float phi1 = (tmp1 + tmp2) * .5f;
      tmp2 = (p[-Y] + p[-X - Y]) * .5f;
      float phi2 = (tmp1 + tmp2) * .5f;
      dx = p[0] - p[-X];
      dy = phi1 - phi2;
      if (g.is3D) {
        tmp2 = (p[Z] + p[-X + Z]) * .5f; phi1 = (tmp1 + tmp2) * .5f;
        tmp2 = (p[-Z] + p[-X - Z]) * .5f; phi2 = (tmp1 + tmp2) * .5f;
        dz = phi1 - phi2;
      }
      normalize3(dx, dy, dz);
      float mx, my, mz; getAtMACX(vel, g, idx, mx, my, mz);
      tx = mx - (dx * mx + dy * my + dz * mz) * dx;
    }
    if (curObs | fIsObstacle(flags[idx - Y])) {
      float dx = 0.f, dy = 0.f, dz = 0.f;
      const float tmp1 = (p[0] + p[-Y]) * .5f;
      float tmp2 = (p[X] + p[X - Y]) * .5f;
      float phi1 = (tmp1 + tmp2) * .5f;
      tmp2 = (p[-X] + p[-X - Y]) * .5f;
      float phi2 = (tmp1 + tmp2) * .5f;
      dx = phi1 - phi2;
      dy = p[0] - p[-Y];
      if (g.is3D) {
        tmp2 = (p[Z] + p[-Y + Z]) * .5f; phi1 = (tmp1 + tmp2) * .5f;
        tmp2 = (p[-Z] + p[-Y - Z]) * .5f; phi2 = (tmp1 + tmp2) * .5f;
        dz = phi1 - phi2;
      }
      normalize3(dx, dy, dz);
      float mx, my, mz; getAtMACY(vel, g, idx, mx, my, mz);
      ty = my - (dx * mx + dy * my + dz * mz) * dy;
    }
    if (g.is3D && (curObs | fIsObstacle(flags[idx - Z]))) {
      float dx = 0.f, dy = 0.f, dz = 0.f;
      const float tmp1 = (p[0] + p[-Z]) * .5f;
      float tmp2 = (p[X] + p[X - Z]) * .5f;
      float phi1 = (tmp1 + tmp2) * .5f;
      tmp2 = (p[-X] + p[-X - Z]) * .5f;
      float phi2 = (tmp1 + tmp2) * .5f;
      dx = phi1 - phi2;
      tmp2 = (p[Y] + p[Y - Z]) * .5f; phi1 = (tmp1 + tmp2) * .5f;
      tmp2 = (p[-Y] + p[-Y - Z]) * .5f; phi2 = (tmp1 + tmp2) * .5f;
      dy = phi1 - phi2;
      dz = p[0] - p[-Z];
      normalize3(dx, dy, dz);
      float mx, my, mz; getAtMACZ(vel, g, idx, mx, my, mz);
      tz = mz - (dx * mx + dy * my + dz * mz) * dz;
    }
  }
  tgt.x[idx] = tx; tgt.y[idx] = ty; tgt.z[idx] = tz;
}

static void parseDesc(const char* desc, int& lo, int& up, const char* fn) {
  lo = up = 0;
  for (const char* c = desc; c && *c; c++) {
    if (*c == 'x') lo |= 1; else if (*c == 'y') lo |= 2; else if (*c == 'z') lo |= 4;
    else if (*c == 'X') up |= 1; else if (*c == 'Y') up |= 2; else if (*c == 'Z') up |= 4;
    else mfErr("invalid character in boundary description string. Only [xyzXYZ] allowed.");
  }
  (void)fn;
}

} // namespace mf

using namespace mf;

extern "C" {

int mf_flags_init_domain(mf_grid* fp, int boundaryWidth, const char* wallIn, const char* openIn,
                         const char* inflowIn, const char* outflowIn, mf_grid* phiWalls) {
  MF_API_BEGIN
  REQ_GRID(fp, G(fp)->isInt(), "flags", "FlagGrid::initDomain");
  Grid* flags = G(fp); Solver* s = flags->s; s->use();
  markFlagsChanged(flags);
  std::string wall = wallIn ? wallIn : "xXyYzZ", open = openIn ? openIn : "", inflow = inflowIn ? inflowIn : "", outflow = outflowIn ? outflowIn : "";
  wall.append("      "); open.append("      "); inflow.append("      "); outflow.append("      ");
  Types6 types; bool set[6] = { false, false, false, false, false, false };
  for (int i = 0; i < 6; i++) types.t[i] = 0;
  int wallMask = 0;
  const char lower[6] = { 'x', 'X', 'y', 'Y', 'z', 'Z' };
  for (int i = 0; i < 6; ++i) {           // grid.cpp:813-882
    for (int side = 0; side < (s->is3D ? 6 : 4); side++) {
      if (set[side]) continue;
      const char c = lower[side];
      if (open[i] == c) { types.t[side] = MF_OPEN; set[side] = true; }
      else if (inflow[i] == c) { types.t[side] = MF_INFLOW; set[side] = true; }
      else if (outflow[i] == c) { types.t[side] = MF_OUTFLOW; set[side] = true; }
      else if (wall[i] == c) { types.t[side] = MF_OBSTACLE; wallMask |= (1 << side); set[side] = true; }
    }
  }
  if (phiWalls) {
    REQ_GRID(phiWalls, G(phiWalls)->isReal(), "phiWalls", "FlagGrid::initDomain");
    MF_LAUNCH_IJK(s, k_init_phi_walls, G(phiWalls)->d, boundaryWidth, wallMask);
  }
  MF_LAUNCH_IJK(s, k_init_domain, flags->di(), boundaryWidth, types);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_flags_fill_grid(mf_grid* fp, int type) {
  MF_API_BEGIN
  REQ_GRID(fp, G(fp)->isInt(), "flags", "FlagGrid::fillGrid");
  Grid* flags = G(fp); Solver* s = flags->s; s->use();
  markFlagsChanged(flags);
  MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH(s, k_fill_grid, s->numSMs * 8, 256, flags->di() + sp.off, type, sp.n); }
  s->rankSync();
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_flags_count_cells(const mf_grid* fp, int flag, int bnd, int* out) {
  MF_API_BEGIN
  REQ_GRID(fp, G(fp)->isInt(), "flags", "FlagGrid::countCells");
  mfAssert(out, "FlagGrid::countCells: null output");
  const Grid* flags = G(fp); Solver* s = flags->s; s->use();
  s->clearShared(s->devFlagsScratch + 1, sizeof(int));
  MF_LAUNCH_IJK(s, k_count_cells, flags->di(), flag, bnd, s->devFlagsScratch + 1);
  MF_LAUNCH_CHECK();
  s->readSmall(s->hostInts, s->devFlagsScratch + 1, (int)sizeof(int));
  *out = s->hostInts[0];
  MF_API_END
}

int mf_flags_set_open_bound(mf_grid* fp, int bWidth, const char* openBound, int type) {
  MF_API_BEGIN
  REQ_GRID(fp, G(fp)->isInt(), "flags", "setOpenBound");
  if (!openBound || !*openBound) return 0;
  Grid* flags = G(fp); Solver* s = flags->s; s->use();
  markFlagsChanged(flags);
  int lo, up; parseDesc(openBound, lo, up, "setOpenBound");
  MF_LAUNCH_IJK(s, k_set_open_bound, flags->di(), bWidth, lo, up, type);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_reset_outflow(mf_grid* fp, mf_grid* phi, mf_grid* real) {
  MF_API_BEGIN
  REQ_GRID(fp, G(fp)->isInt(), "flags", "resetOutflow");
  Grid* flags = G(fp); Solver* s = flags->s; s->use();
  if (phi) REQ_GRID(phi, G(phi)->isReal(), "phi", "resetOutflow");
  if (real) REQ_GRID(real, G(real)->isReal(), "real", "resetOutflow");
  { const int o = flags->outflowState; markFlagsChanged(flags); flags->outflowState = o; }   // fluid / empty bits change, the outflow bit does not
  MF_RANKS(s, r) { const Span sp = span(s, r);
    MF_LAUNCH(s, k_reset_outflow, s->numSMs * 8, 256, flags->di() + sp.off, phi ? G(phi)->d + sp.off : nullptr, real ? G(real)->d + sp.off : nullptr, sp.n); }
  s->rankSync();
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_apply_box_to_real(mf_grid* gp, float x0, float y0, float z0, float x1, float y1, float z1, float value) {
  MF_API_BEGIN
  REQ_GRID(gp, G(gp)->isReal(), "grid", "Shape::applyToGrid");
  Grid* gr = G(gp); Solver* s = gr->s; s->use();
  MF_LAUNCH_IJK(s, k_apply_box<float>, gr->d, x0, y0, z0, x1, y1, z1, value);
  MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_apply_box_to_flags(mf_grid* gp, float x0, float y0, float z0, float x1, float y1, float z1, int value) {
  MF_API_BEGIN
  REQ_GRID(gp, G(gp)->isInt(), "grid", "Shape::applyToGrid");
  Grid* gr = G(gp); Solver* s = gr->s; s->use();
  markFlagsChanged(gr);
  MF_LAUNCH_IJK(s, k_apply_box<int>, gr->di(), x0, y0, z0, x1, y1, z1, value);
  MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_apply_cylinder_to_real(mf_grid* gp, float cx, float cy, float cz, float r, float zx, float zy, float zz, float value) {
  MF_API_BEGIN
  REQ_GRID(gp, G(gp)->isReal(), "grid", "Shape::applyToGrid");
  Grid* gr = G(gp); Solver* s = gr->s; s->use();
  // Cylinder ctor: mZDir = z; mZ = normalize(mZDir)  (shapes.cpp:316-322, vectorbase.h:421-434)
  float zlen;
  {
    const float l = zx * zx + zy * zy + zz * zz;
    if (std::fabs((double)l - 1.) < (double)(1e-6f * 1e-6f)) zlen = 1.f;
    else if (l > 1e-6f * 1e-6f) { zlen = std::sqrt(l); const float inv = (float)(1. / (double)zlen); zx *= inv; zy *= inv; zz *= inv; }
    else { zx = zy = zz = 0.f; zlen = 0.f; }
  }
  MF_LAUNCH_IJK(s, k_apply_cylinder, gr->d, cx, cy, cz, r, zx, zy, zz, zlen, value);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_add_buoyancy(const mf_grid* fp, const mf_grid* dp, mf_grid* vp, float gx, float gy, float gz, float coefficient, int scale) {
  MF_API_BEGIN
  REQ_GRID(fp, G(fp)->isInt(), "flags", "addBuoyancy");
  REQ_GRID(dp, G(dp)->isReal(), "density", "addBuoyancy");
  REQ_GRID(vp, G(vp)->isMAC(), "vel", "addBuoyancy");
  const Grid* flags = G(fp); Solver* s = flags->s; s->use();
  check_same_solver(flags, G(dp), "addBuoyancy"); check_same_solver(flags, G(vp), "addBuoyancy");
  // Vec3 f = -gravity * dt / gridScale * coefficient  (extforces.cpp:85-86), all in float
  const float gridScale = scale ? s->dx : 1.0f;
  const float dt = s->dt;
  const float fx = ((-gx * dt) / gridScale) * coefficient;
  const float fy = ((-gy * dt) / gridScale) * coefficient;
  const float fz = ((-gz * dt) / gridScale) * coefficient;
  MF_LAUNCH_IJK_W_PDL(s, k_add_buoyancy, flags->di(), G(dp)->d, v3(G(vp)), fx, fy, fz);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_add_gravity(const mf_grid* fp, mf_grid* vp, float gx, float gy, float gz, const mf_grid* exclude, int scale) {
  MF_API_BEGIN
  REQ_GRID(fp, G(fp)->isInt(), "flags", "addGravity");
  REQ_GRID(vp, G(vp)->isMAC(), "vel", "addGravity");
  if (exclude) REQ_GRID(exclude, G(exclude)->isReal(), "exclude", "addGravity");
  const Grid* flags = G(fp); Solver* s = flags->s; s->use();
  const float gridScale = scale ? s->dx : 1.0f;
  const float dt = s->dt;
  const float fx = (gx * dt) / gridScale, fy = (gy * dt) / gridScale, fz = (gz * dt) / gridScale;   // extforces.cpp:63-64
  MF_LAUNCH_IJK(s, k_apply_force, flags->di(), v3(G(vp)), fx, fy, fz, exclude ? G(exclude)->d : nullptr, 1);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_set_inflow_bcs(mf_grid* vp, const char* dir, float vx, float vy, float vz) {
  MF_API_BEGIN
  REQ_GRID(vp, G(vp)->isMAC(), "vel", "setInflowBcs");
  Grid* vel = G(vp); Solver* s = vel->s; s->use();
  const int size[3] = { s->nx, s->ny, s->nz };
  for (const char* c = dir; c && *c; c++) {
    if (*c >= 'x' && *c <= 'z') { MF_LAUNCH_IJK(s, k_set_inflow, v3(vel), *c - 'x', 0, vx, vy, vz); }
    else if (*c >= 'X' && *c <= 'Z') { const int d = *c - 'X'; MF_LAUNCH_IJK(s, k_set_inflow, v3(vel), d, size[d] - 1, vx, vy, vz); }
    else mfErr("invalid character in direction string. Only [xyzXYZ] allowed.");
  }
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_set_wall_bcs(const mf_grid* fp, mf_grid* vp, const mf_grid* obvel, const mf_grid* fractions, const mf_grid* phiObs, int boundaryWidth) {
  MF_API_BEGIN
  REQ_GRID(fp, G(fp)->isInt(), "flags", "setWallBcs");
  REQ_GRID(vp, G(vp)->isMAC(), "vel", "setWallBcs");
  const Grid* flags = G(fp); Grid* vel = G(vp); Solver* s = flags->s; s->use();
  check_same_solver(flags, vel, "setWallBcs");
  if (obvel) REQ_GRID(obvel, G(obvel)->isMAC(), "obvel", "setWallBcs");
  if (fractions) REQ_GRID(fractions, G(fractions)->isMAC(), "fractions", "setWallBcs");
  if (phiObs) REQ_GRID(phiObs, G(phiObs)->isReal(), "phiObs", "setWallBcs");
  if (!phiObs || !fractions) {
    CV3 ov; ov.x = ov.y = ov.z = nullptr;
    if (obvel) ov = cv3(G(obvel));
    const int split = (s->is3D && s->nx % 4 == 0 && !getenv("MF_WALLBCS_NOSPLIT")) ? 1 : 0;
    if (split) {
      const dim3 eb(2, 64, 2);
      MF_RANKS(s, r) { const Geo g = geo(s, r);
        if (g.k1 > g.k0) { MF_LAUNCH(s, k_set_wall_bcs_edge<4>, dim3(1, cdiv(g.ny, 64), cdiv(g.k1 - g.k0, 2)), eb, g, flags->di(), v3(vel), ov, obvel ? 1 : 0); } }
    }
    MF_LAUNCH_IJK_W_PDL(s, k_set_wall_bcs, flags->di(), v3(vel), ov, obvel ? 1 : 0, split);
  } else {
    mfAssert(s->nx >= 3 && s->ny >= 3, "setWallBcs: grid too small");
    float* tmp = s->getVec();
    Grid tg = *vel; tg.d = tmp;
    MF_LAUNCH_IJK(s, k_set_wall_bcs_frac, flags->di(), cv3(vel), v3(&tg), G(phiObs)->d);
    float* old = vel->d; vel->d = tmp;     // vel.swap(tmpvel)
    s->putVec(old);
  }
  (void)boundaryWidth;
  MF_LAUNCH_CHECK();
  MF_API_END
}

} // extern "C"
