// mg_kernels.cuh -- device kernels of the multigrid preconditioner (reference: source/multigrid.cpp).
// Layout: level 0 keeps the 7-point matrix as four planes A[d*n + v] (d = 0: diagonal, 1..3: +x,+y,+z coupling,
// i.e. the reference's mA[0][v*4+d], multigrid.cpp:351-358); levels > 0 keep the symmetric 27-point operator as
// 14 (5 in 2-D) planes A[e*n + v] (reference: mA[l][v*14+e], :208-218).  type[] holds GridMg::VertexType.
#pragma once
#include "common.cuh"
#include "reduce.cuh"

namespace mf {

enum : unsigned char { vtInactive = 0, vtActive = 1, vtActiveTrivial = 2, vtRemoved = 3, vtZero = 4, vtFree = 5 };

struct LvlD {
  int sx, sy, sz;
  int n;
  float* A; float* x; float* b; float* r;
  unsigned char* type;
};

struct MgDims { int is3D; int dim; int SS; int SS0; int zmin; int zmax; };   // mStencilSize, mStencilSize0, mStencilMin/Max.z

#define MG_V_FROM_THREAD(L) \
  const int v = blockIdx.x * blockDim.x + threadIdx.x; \
  if (v >= (L).n) return; \
  const int vx = v % (L).sx, vy = (v % ((L).sx * (L).sy)) / (L).sx, vz = v / ((L).sx * (L).sy);

__device__ __forceinline__ bool mgInGrid(const LvlD& L, int x, int y, int z) {
  return x >= 0 && y >= 0 && z >= 0 && x < L.sx && y < L.sy && z < L.sz;
}
__device__ __forceinline__ int mgLin(const LvlD& L, int x, int y, int z) { return x + y * L.sx + z * L.sx * L.sy; }

// knCopyA + knActivateVertices + analyzeStencil (multigrid.cpp:321-384)
__global__ void __launch_bounds__(256) k_mg_copy_activate(LvlD L, MgDims D, const float* __restrict__ A0, const float* __restrict__ Ai,
                                                          const float* __restrict__ Aj, const float* __restrict__ Ak, float trivialScale,
                                                          int* __restrict__ found /* [0]=nonZeroStencilSum, [1]=trivial */) {
  MG_V_FROM_THREAD(L)
  const int n = L.n;
  float a[7];
  a[0] = A0[v]; a[1] = Ai[v]; a[2] = Aj[v]; a[3] = D.is3D ? Ak[v] : 0.f;
  a[4] = vx != 0 ? Ai[v - 1] : 0.f;
  a[5] = vy != 0 ? Aj[v - L.sx] : 0.f;
  a[6] = (vz != 0 && D.is3D) ? Ak[v - L.sx * L.sy] : 0.f;
  unsigned char t = vtInactive;
  float diag = a[0];
  if (a[0] != 0.f) {
    t = vtActive;
    float stencilMax = 0.f, stencilSum = 0.f;
    for (int q = 0; q < 7; q++) { stencilSum += a[q]; stencilMax = fmaxf(stencilMax, fabsf(a[q])); }
    if (fabsf(stencilSum / stencilMax) > 1E-6f) found[0] = 1;
    const bool trivial = a[0] == 1.f && a[1] == 0.f && a[2] == 0.f && a[3] == 0.f && a[4] == 0.f && a[5] == 0.f && a[6] == 0.f;
    if (trivial) { t = vtActiveTrivial; diag = a[0] * trivialScale; found[1] = 1; }
  }
  L.type[v] = t;
  L.A[v] = diag; L.A[n + v] = a[1]; L.A[2 * n + v] = a[2];
  if (D.is3D) L.A[3 * n + v] = a[3];
}

// knSetRhs (multigrid.cpp:417-424), only launched when trivial equations exist
__global__ void __launch_bounds__(256) k_mg_set_rhs(LvlD L, const float* __restrict__ rhs, float* __restrict__ b, float trivialScale) {
  MG_V_FROM_THREAD(L)
  float r = rhs[v];
  if (L.type[v] == vtActiveTrivial) r *= trivialScale;
  b[v] = r;
}

__global__ void k_mg_fill_type(unsigned char* t, unsigned char val, int n) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < n) t[v] = val;
}

// one sweep of the order-independent part of genCoarseGrid (multigrid.cpp:520-578): a fine vertex that is still in the
// heap with exactly one free interpolation vertex selects it ('zero').  flags[0] = changed, flags[1] = some vertex still
// has >= 2 free interpolation vertices (the serial algorithm would have to pop it: order-dependent -> host fallback).
__global__ void __launch_bounds__(256) k_mg_coarsen_sweep(LvlD F, LvlD C, int* __restrict__ flags) {
  MG_V_FROM_THREAD(F)
  if (F.type[v] == vtInactive) return;
  int cnt = 0, last = -1;
  for (int z = vz / 2; z <= (vz + 1) / 2; z++)
    for (int y = vy / 2; y <= (vy + 1) / 2; y++)
      for (int x = vx / 2; x <= (vx + 1) / 2; x++) {
        const int i = mgLin(C, x, y, z);
        if (((volatile unsigned char*)C.type)[i] == vtFree) { cnt++; last = i; }
      }
  if (cnt == 1) { C.type[last] = vtZero; flags[0] = 1; }
  else if (cnt >= 2) flags[1] = 1;
}
// knActivateCoarseVertices (multigrid.cpp:507-516)
__global__ void k_mg_coarsen_finalize(unsigned char* t, int n) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n) return;
  unsigned char c = t[v];
  if (c == vtFree) c = vtRemoved;
  if (c == vtZero) c = vtActive;
  if (c == vtRemoved) c = vtInactive;
  t[v] = c;
}

// precomputed coarsening paths for level 1 (multigrid.cpp:286-318)
struct MgPath { signed char Nx, Ny, Nz, Ux, Uy, Uz, Wx, Wy, Wz; signed char sc, sf, inU; float rw, iw; };

// knGenCoarseGridOperator, l == 1 (multigrid.cpp:594-614)
__global__ void __launch_bounds__(128) k_mg_galerkin1(LvlD F, LvlD C, MgDims D, const MgPath* __restrict__ paths, int npaths) {
  MG_V_FROM_THREAD(C)
  if (C.type[v] == vtInactive) { for (int e = 0; e < D.SS; e++) C.A[(long long)e * C.n + v] = 0.f; return; }
  float acc[14];
#pragma unroll
  for (int e = 0; e < 14; e++) acc[e] = 0.f;
  for (int q = 0; q < npaths; q++) {
    const MgPath p = paths[q];
    const int Nx = vx + p.Nx, Ny = vy + p.Ny, Nz = vz + p.Nz;
    if (!mgInGrid(C, Nx, Ny, Nz) || C.type[mgLin(C, Nx, Ny, Nz)] == vtInactive) continue;
    const int Ux = vx * 2 + p.Ux, Uy = vy * 2 + p.Uy, Uz = vz * 2 + p.Uz;
    if (!mgInGrid(F, Ux, Uy, Uz)) continue;
    const int u = mgLin(F, Ux, Uy, Uz);
    if (F.type[u] == vtInactive) continue;
    const int Wx = vx * 2 + p.Wx, Wy = vy * 2 + p.Wy, Wz = vz * 2 + p.Wz;
    if (!mgInGrid(F, Wx, Wy, Wz)) continue;
    const int w = mgLin(F, Wx, Wy, Wz);
    if (F.type[w] == vtInactive) continue;
    const float a = F.A[(long long)p.sf * F.n + (p.inU ? u : w)];
    const float add = p.rw * a * p.iw;
    // static indexing keeps acc[] in registers
#pragma unroll
    for (int e = 0; e < 14; e++) if (e == p.sc) acc[e] += add;
  }
  for (int e = 0; e < D.SS; e++) C.A[(long long)e * C.n + v] = acc[e];
}

// knGenCoarseGridOperator, l > 1 (multigrid.cpp:615-656); F has the 27-point layout
__global__ void __launch_bounds__(128) k_mg_galerkin(LvlD F, LvlD C, MgDims D) {
  MG_V_FROM_THREAD(C)
  if (C.type[v] == vtInactive) { for (int e = 0; e < D.SS; e++) C.A[(long long)e * C.n + v] = 0.f; return; }
  float acc[14];
#pragma unroll
  for (int e = 0; e < 14; e++) acc[e] = 0.f;
  const int SS = D.SS;
  const int smz = D.is3D ? 1 : 0;
  for (int Uz = max(0, vz * 2 - 1); Uz <= min(F.sz - 1, vz * 2 + 1); Uz++)
  for (int Uy = max(0, vy * 2 - 1); Uy <= min(F.sy - 1, vy * 2 + 1); Uy++)
  for (int Ux = max(0, vx * 2 - 1); Ux <= min(F.sx - 1, vx * 2 + 1); Ux++) {
    const int u = mgLin(F, Ux, Uy, Uz);
    if (F.type[u] == vtInactive) continue;
    const float rw = 1.0f / (float)(1 << ((Ux % 2) + (Uy % 2) + (Uz % 2)));
    // (U-1)/2 uses C++ truncating division, as in the reference
    for (int Nz = (Uz - 1) / 2; Nz <= min(C.sz - 1, (Uz + 2) / 2); Nz++)
    for (int Ny = (Uy - 1) / 2; Ny <= min(C.sy - 1, (Uy + 2) / 2); Ny++)
    for (int Nx = (Ux - 1) / 2; Nx <= min(C.sx - 1, (Ux + 2) / 2); Nx++) {
      const int nn = mgLin(C, Nx, Ny, Nz);
      if (C.type[nn] == vtInactive) continue;
      const int sc = (Nx - vx + 1) + 3 * (Ny - vy + 1) + 9 * (Nz - vz + smz);
      if (sc < SS - 1) continue;
      const int e = sc - SS + 1;
      for (int Wz = max(0, max(Uz - 1, Nz * 2 - 1)); Wz <= min(F.sz - 1, min(Uz + 1, Nz * 2 + 1)); Wz++)
      for (int Wy = max(0, max(Uy - 1, Ny * 2 - 1)); Wy <= min(F.sy - 1, min(Uy + 1, Ny * 2 + 1)); Wy++)
      for (int Wx = max(0, max(Ux - 1, Nx * 2 - 1)); Wx <= min(F.sx - 1, min(Ux + 1, Nx * 2 + 1)); Wx++) {
        const int w = mgLin(F, Wx, Wy, Wz);
        if (F.type[w] == vtInactive) continue;
        const int sf = (Wx - Ux + 1) + 3 * (Wy - Uy + 1) + 9 * (Wz - Uz + smz);
        const float iw = 1.0f / (float)(1 << ((Wx % 2) + (Wy % 2) + (Wz % 2)));
        float a;
        if (sf < SS) a = F.A[(long long)(SS - 1 - sf) * F.n + w];
        else         a = F.A[(long long)(sf - SS + 1) * F.n + u];
        const float add = rw * a * iw;
#pragma unroll
        for (int q = 0; q < 14; q++) if (q == e) acc[q] += add;
      }
    }
  }
  for (int e = 0; e < SS; e++) C.A[(long long)e * C.n + v] = acc[e];
}

// ---- level 0 (7-point) -----------------------------------------------------------------------------
// b - sum_offdiag(A x) in the reference's order (-x,+x,-y,+y,-z,+z; multigrid.cpp:686-689, :750-753)
__device__ __forceinline__ float mgOffdiag0(const LvlD& L, const MgDims& D, int v, int vx, int vy, int vz, float sum) {
  const int n = L.n, Y = L.sx, Z = L.sx * L.sy;
  if (vx > 0)        sum -= L.A[n + v - 1] * L.x[v - 1];
  if (vx < L.sx - 1) sum -= L.A[n + v] * L.x[v + 1];
  if (vy > 0)        sum -= L.A[2 * n + v - Y] * L.x[v - Y];
  if (vy < L.sy - 1) sum -= L.A[2 * n + v] * L.x[v + Y];
  if (D.is3D) {
    if (vz > 0)        sum -= L.A[3 * (long long)n + v - Z] * L.x[v - Z];
    if (vz < L.sz - 1) sum -= L.A[3 * (long long)n + v] * L.x[v + Z];
  }
  return sum;
}
// knSmoothColor, l == 0: all vertices with (x+y+z)&1 == color (the reference's two colour groups, :721,:724)
__global__ void __launch_bounds__(256) k_mg_smooth0(LvlD L, MgDims D, int color) {
  const int hx = (L.sx + 1) / 2;
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= (long long)hx * L.sy * L.sz) return;
  const int ii = (int)(t % hx), vy = (int)((t / hx) % L.sy), vz = (int)(t / ((long long)hx * L.sy));
  const int vx = 2 * ii + ((vy + vz + color) & 1);
  if (vx >= L.sx) return;
  const int v = mgLin(L, vx, vy, vz);
  if (L.type[v] == vtInactive) return;
  const float sum = mgOffdiag0(L, D, v, vx, vy, vz, L.b[v]);
  L.x[v] = sum / L.A[v];
}
// knCalcResidual, l == 0
__global__ void __launch_bounds__(256) k_mg_residual0(LvlD L, MgDims D) {
  MG_V_FROM_THREAD(L)
  if (L.type[v] == vtInactive) return;
  float sum = mgOffdiag0(L, D, v, vx, vy, vz, L.b[v]);
  sum -= L.A[v] * L.x[v];
  L.r[v] = sum;
}

// ---- levels > 0 (27-point, 14 stored) -----------------------------------------------------------------
template <bool WITH_CENTER>
__device__ __forceinline__ float mgStencilSub(const LvlD& L, const MgDims& D, int v, int vx, int vy, int vz, float sum) {
  const int SS = D.SS;
  int s = 0;
  for (int dz = D.zmin; dz <= D.zmax; dz++)
    for (int dy = -1; dy <= 1; dy++)
      for (int dx = -1; dx <= 1; dx++, s++) {
        if (!WITH_CENTER && s == SS - 1) continue;
        const int Nx = vx + dx, Ny = vy + dy, Nz = vz + dz;
        if (!mgInGrid(L, Nx, Ny, Nz)) continue;
        const int nn = mgLin(L, Nx, Ny, Nz);
        if (L.type[nn] == vtInactive) continue;
        if (s < SS) sum -= L.A[(long long)(SS - 1 - s) * L.n + nn] * L.x[nn];
        else        sum -= L.A[(long long)(s - SS + 1) * L.n + v] * L.x[nn];
      }
  return sum;
}
// knSmoothColor, l > 0: one of the 8 (4) colours = corner offset of the 2x2(x2) blocks (multigrid.cpp:668-711)
__global__ void __launch_bounds__(128) k_mg_smooth(LvlD L, MgDims D, int ox, int oy, int oz) {
  const int bx = (L.sx + 1) / 2, by = (L.sy + 1) / 2, bz = (L.sz + 1) / 2;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= bx * by * bz) return;
  const int vx = (t % bx) * 2 + ox, vy = ((t % (bx * by)) / bx) * 2 + oy, vz = (t / (bx * by)) * 2 + oz;
  if (!mgInGrid(L, vx, vy, vz)) return;
  const int v = mgLin(L, vx, vy, vz);
  if (L.type[v] == vtInactive) return;
  const float sum = mgStencilSub<false>(L, D, v, vx, vy, vz, L.b[v]);
  L.x[v] = sum / L.A[v];
}
__global__ void __launch_bounds__(128) k_mg_residual(LvlD L, MgDims D) {
  MG_V_FROM_THREAD(L)
  if (L.type[v] == vtInactive) return;
  L.r[v] = mgStencilSub<true>(L, D, v, vx, vy, vz, L.b[v]);
}

// knRestrict (multigrid.cpp:904-927): b_coarse = R r_fine ; also x_coarse = 0 (knSet, :472)
__global__ void __launch_bounds__(128) k_mg_restrict(LvlD F, LvlD C) {
  MG_V_FROM_THREAD(C)
  C.x[v] = 0.f;
  if (C.type[v] == vtInactive) return;
  float sum = 0.f;
  for (int z = max(0, vz * 2 - 1); z <= min(F.sz - 1, vz * 2 + 1); z++)
    for (int y = max(0, vy * 2 - 1); y <= min(F.sy - 1, vy * 2 + 1); y++)
      for (int x = max(0, vx * 2 - 1); x <= min(F.sx - 1, vx * 2 + 1); x++) {
        const int r = mgLin(F, x, y, z);
        if (F.type[r] == vtInactive) continue;
        const float rw = 1.0f / (float)(1 << ((x % 2) + (y % 2) + (z % 2)));
        sum += rw * F.r[r];
      }
  C.b[v] = sum;
}
// knInterpolate + knAddAssign (multigrid.cpp:934-954, :445-446): x_fine += P x_coarse
__global__ void __launch_bounds__(256) k_mg_interp_add(LvlD F, LvlD C) {
  MG_V_FROM_THREAD(F)
  if (F.type[v] == vtInactive) return;
  float sum = 0.f;
  for (int z = vz / 2; z <= (vz + 1) / 2; z++)
    for (int y = vy / 2; y <= (vy + 1) / 2; y++)
      for (int x = vx / 2; x <= (vx + 1) / 2; x++) {
        const int i = mgLin(C, x, y, z);
        if (C.type[i] != vtInactive) sum += C.x[i];
      }
  const float iw = 1.0f / (float)(1 << ((vx % 2) + (vy % 2) + (vz % 2)));
  F.x[v] += iw * sum;
}

} // namespace mf
