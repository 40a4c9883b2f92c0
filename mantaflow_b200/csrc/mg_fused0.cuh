// mg_fused0.cuh -- fused level-0 legs of the V(1,1)-cycle (3-D).  Level 0 is the only level whose vectors do not fit the L2,
// so every separate sweep is a full HBM round trip; the reference's sequence (multigrid.cpp:448-504)
//     x = 0; smooth(colour 0); smooth(colour 1); residual; restrict            ... coarse levels ...
//     interpolate + add; smooth(colour 1); smooth(colour 0)
// is regrouped so that intermediate values stay on chip.  Every value is computed by the same float expression as in the unfused
// kernels (mg_kernels.cuh), only redundantly in the tile halos: results are bit-identical.  (A fused residual+restriction tile
// kernel was measured too: 209 us against 95 + 58 us for the two separate sweeps at 256^3 -- instruction-bound -- and dropped.)
//   k_mg_presmooth0   both colours of the first sweep from x = 0: a first-colour value is rhs/diag, so a second-colour vertex
//                     recomputes its six neighbours' values from b and the operator row instead of reading x.   b,row -> x
//   k_mg_up0          interpolation-add + both colours of the post sweep on a tile with a 2-cell halo.          x,coarse x,b,row -> x'
#pragma once
#include "mg_kernels.cuh"

namespace mf {

// b - sum_offdiag(A x) with the x values taken from a (shared-memory) tile: xs[si] is vertex v, strides 1 / SY / SZ
// EDGE = false: the CTA's whole working region lies strictly inside the grid, so the six boundary guards (and the in-grid tests
// of the callers) are compiled out -- true for all but the outermost shell of tiles
template <class OP, bool EDGE>
__device__ __forceinline__ float mgOffdiag0T(const OP& op, const typename OP::Row& row, const LvlD& L, const float* xs, int si, int SY, int SZ,
                                             int v, int vx, int vy, int vz, float sum) {
  const int Y = L.sx, Z = L.sx * L.sy;
  if (!EDGE || vx > 0)        sum = op.template sub<0>(row, v, v - 1, sum, xs[si - 1]);
  if (!EDGE || vx < L.sx - 1) sum = op.template sub<1>(row, v, v + 1, sum, xs[si + 1]);
  if (!EDGE || vy > 0)        sum = op.template sub<2>(row, v, v - Y, sum, xs[si - SY]);
  if (!EDGE || vy < L.sy - 1) sum = op.template sub<3>(row, v, v + Y, sum, xs[si + SY]);
  if (!EDGE || vz > 0)        sum = op.template sub<4>(row, v, v - Z, sum, xs[si - SZ]);
  if (!EDGE || vz < L.sz - 1) sum = op.template sub<5>(row, v, v + Z, sum, xs[si + SZ]);
  return sum;
}

// value a first-colour vertex n gets in the first sweep from x = 0 (k_mg_smooth0<XZERO>)
template <class OP>
__device__ __forceinline__ float mgFirstValue(const OP& op, const float* __restrict__ b, int n) {
  const typename OP::Row r = op.row(n);
  if (op.inactive(r)) return 0.f;
  return op.rhs(r, b[n]) / op.diag(r, n);
}

// smooth(colour firstColor) then smooth(1 - firstColor), starting from x = 0.  One warp per (y,z) row of [z0,z1); a lane owns
// a PAIR of x-neighbours (one cell of each colour, so no lane idles on a colour test): it computes the first-colour value of
// its pair, takes the other x-neighbour's from the adjacent lane by shuffle and recomputes the four y/z neighbours' values.
template <class OP>
__global__ void __launch_bounds__(256) k_mg_presmooth0(LvlD L, float trivialScale, int firstColor, int z0, int z1) {
  MF_PDL_PROLOGUE
  const OP op(L, trivialScale);
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5) + z0 * L.sy;
  if (row >= L.sy * z1) return;
  const int lane = threadIdx.x & 31;
  const int vy = row % L.sy, vz = row / L.sy;
  const int base = row * L.sx;
  const int Y = L.sx, Z = L.sx * L.sy;
  const float* __restrict__ b = L.b;
  const int par = (vy + vz + firstColor) & 1;             // x parity of the first-colour cells of this row
  for (int x0 = 0; x0 < L.sx; x0 += 64) {
    const int xf = x0 + 2 * lane + par, xs = x0 + 2 * lane + 1 - par;     // first / second colour cell of the pair
    const bool inF = xf < L.sx, inS = xs < L.sx;
    const float fv = inF ? mgFirstValue(op, b, base + xf) : 0.f;
    if (inF) L.x[base + xf] = fv;
    // the second-colour cell's x-neighbours: one is the pair's own first-colour cell, the other the adjacent lane's
    const float fvUp = __shfl_up_sync(0xffffffffu, fv, 1), fvDn = __shfl_down_sync(0xffffffffu, fv, 1);
    if (!inS) continue;
    const int v = base + xs;
    const typename OP::Row r = op.row(v);
    float out = 0.f;
    if (!op.inactive(r)) {
      float xm, xp;                                       // first-sweep values at xs-1, xs+1
      if (par == 0) { xm = fv; xp = (lane < 31) ? fvDn : ((xs + 1 < L.sx) ? mgFirstValue(op, b, v + 1) : 0.f); }
      else          { xp = fv; xm = (lane > 0) ? fvUp : ((xs > 0) ? mgFirstValue(op, b, v - 1) : 0.f); }
      float sum = op.rhs(r, b[v]);
      if (xs > 0)        sum = op.template sub<0>(r, v, v - 1, sum, xm);
      if (xs < L.sx - 1) sum = op.template sub<1>(r, v, v + 1, sum, xp);
      if (vy > 0)        sum = op.template sub<2>(r, v, v - Y, sum, mgFirstValue(op, b, v - Y));
      if (vy < L.sy - 1) sum = op.template sub<3>(r, v, v + Y, sum, mgFirstValue(op, b, v + Y));
      if (vz > 0)        sum = op.template sub<4>(r, v, v - Z, sum, mgFirstValue(op, b, v - Z));
      if (vz < L.sz - 1) sum = op.template sub<5>(r, v, v + Z, sum, mgFirstValue(op, b, v + Z));
      out = sum / op.diag(r, v);
    }
    L.x[v] = out;
  }
}

// ---- interpolation-add + post sweep ---------------------------------------------------------------------------------
// CTA (256 threads = 8 warps) owns a fine tile UP_X x UP_Y x UP_Z and works on it with a 2-cell halo in shared memory.
// Work is laid out so that index arithmetic is additions only: a warp walks (y,z) rows, its lanes cover the 64 own columns of
// the row (two per lane when every cell is touched, one per lane in a colour pass: lane <-> cells of ONE colour, no divergence);
// the few halo columns left and right are handled by a short flat loop.
constexpr int UP_X = 64, UP_Y = 8, UP_Z = 8;                                    // fine tile owned by a CTA
constexpr int UX_X = UP_X + 4, UX_Y = UP_Y + 4, UX_Z = UP_Z + 4;                 // with the 2-cell halo
constexpr int UC_X = UP_X / 2 + 3, UC_Y = UP_Y / 2 + 3, UC_Z = UP_Z / 2 + 3;     // coarse vertices under it

__device__ __forceinline__ float mgPow2Neg(int k) { return __int_as_float((127 - k) << 23); }   // exactly 1 / 2^k

template <class OP, bool EDGE>
__device__ __forceinline__ void up0Body(const LvlD& F, const LvlD& C, float trivialScale, const float* __restrict__ xin, float* __restrict__ xout,
                                        int firstColor, int z0, int z1, float* sX, float* sC) {
  const OP op(F, trivialScale);
  const int ox = blockIdx.x * UP_X, oy = blockIdx.y * UP_Y, oz = z0 + blockIdx.z * UP_Z;
  const int hx = ox - 2, hy = oy - 2, hz = oz - 2;                               // origin of the halo region
  const int CX0 = hx >> 1, CY0 = hy >> 1, CZ0 = hz >> 1;                          // floor: -1 at the low border (padding entry)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* __restrict__ b = F.b;
  for (int t = threadIdx.x; t < UC_X * UC_Y * UC_Z; t += 256) {
    const int X = CX0 + t % UC_X, Y = CY0 + (t / UC_X) % UC_Y, Z = CZ0 + t / (UC_X * UC_Y);
    // coarse x is 0 at inactive vertices and in the padding; beyond the padding nothing is ever referenced
    sC[t] = (X >= -1 && Y >= -1 && Z >= -1 && X <= C.sx && Y <= C.sy && Z <= C.sz) ? C.x[mgIdx(C, X, Y, Z)] : 0.f;
  }
  __syncthreads();
  // ---- x + P x_coarse on the halo region (knInterpolate + knAddAssign) -> sX
  auto interp = [&](int lx, int ly, int lz) {
    const int vx = hx + lx, vy = hy + ly, vz = hz + lz;
    float val = 0.f;
    if (!EDGE || mgInGrid(F, vx, vy, vz)) {
      const int v = vx + F.sx * (vy + F.sy * vz);
      val = xin[v];
      if (!op.inactive(op.row(v))) {
        const int px = vx & 1, py = vy & 1, pz = vz & 1;
        const float* c = sC + ((vx >> 1) - CX0) + UC_X * (((vy >> 1) - CY0) + UC_Y * ((vz >> 1) - CZ0));
        // the reference's loop nest z, y, x over [v/2, (v+1)/2]; absent terms add +0 (the running sum is never -0)
        float sum = 0.f;
        sum += c[0];
        sum += px ? c[1] : 0.f;
        sum += py ? c[UC_X] : 0.f;
        sum += (px & py) ? c[UC_X + 1] : 0.f;
        sum += pz ? c[UC_X * UC_Y] : 0.f;
        sum += (pz & px) ? c[UC_X * UC_Y + 1] : 0.f;
        sum += (pz & py) ? c[UC_X * UC_Y + UC_X] : 0.f;
        sum += (pz & py & px) ? c[UC_X * UC_Y + UC_X + 1] : 0.f;
        val += mgPow2Neg(px + py + pz) * sum;
      }
    }
    sX[lx + UX_X * (ly + UX_Y * lz)] = val;
  };
  for (int row = warp; row < UX_Y * UX_Z; row += 8) {
    const int ly = row % UX_Y, lz = row / UX_Y;
    interp(2 + lane, ly, lz);
    interp(2 + 32 + lane, ly, lz);
  }
  for (int t = threadIdx.x; t < 4 * UX_Y * UX_Z; t += 256) {                     // halo columns lx = 0, 1, 66, 67
    const int q = t & 3, row = t >> 2;
    interp(q < 2 ? q : UP_X + q, row % UX_Y, row / UX_Y);
  }
  __syncthreads();
  // ---- two colour sweeps: colour firstColor on the tile + 1 cell, then the other colour on the tile itself
  auto relax = [&](int lx, int ly, int lz) {                                     // coordinates in the halo region
    const int vx = hx + lx, vy = hy + ly, vz = hz + lz;
    if (EDGE && !mgInGrid(F, vx, vy, vz)) return;
    const int v = vx + F.sx * (vy + F.sy * vz);
    const typename OP::Row r = op.row(v);
    if (op.inactive(r)) return;
    const int si = lx + UX_X * (ly + UX_Y * lz);
    const float sum = mgOffdiag0T<OP, EDGE>(op, r, F, sX, si, UX_X, UX_X * UX_Y, v, vx, vy, vz, op.rhs(r, b[v]));
    sX[si] = sum / op.diag(r, v);
  };
  // pass 0: rows of the tile + 1 (ly, lz in [1, UP+2]), own columns lx in [2, 65] by lane, columns 1 and 66 by the flat loop
  for (int row = warp; row < (UP_Y + 2) * (UP_Z + 2); row += 8) {
    const int ly = 1 + row % (UP_Y + 2), lz = 1 + row / (UP_Y + 2);
    const int par = (hx + hy + ly + hz + lz + firstColor) & 1;                   // parity lx must have: (vx+vy+vz+color) even
    relax(2 + 2 * lane + par, ly, lz);
  }
  for (int t = threadIdx.x; t < 2 * (UP_Y + 2) * (UP_Z + 2); t += 256) {
    const int lx = (t & 1) ? UP_X + 2 : 1, row = t >> 1;
    const int ly = 1 + row % (UP_Y + 2), lz = 1 + row / (UP_Y + 2);
    if (((hx + lx + hy + ly + hz + lz + firstColor) & 1) == 0) relax(lx, ly, lz);
  }
  __syncthreads();
  // pass 1: the other colour on the own tile
  for (int row = warp; row < UP_Y * UP_Z; row += 8) {
    const int ly = 2 + row % UP_Y, lz = 2 + row / UP_Y;
    const int par = (hx + hy + ly + hz + lz + 1 - firstColor) & 1;
    relax(2 + 2 * lane + par, ly, lz);
  }
  __syncthreads();
  for (int row = warp; row < UP_Y * UP_Z; row += 8) {
    const int ly = row % UP_Y, lz = row / UP_Y;
    const int vy = oy + ly, vz = oz + lz;
    if (vy >= F.sy || vz >= F.sz || vz >= z1) continue;
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int vx = ox + lane + 32 * h;
      if (vx < F.sx) xout[vx + F.sx * (vy + F.sy * vz)] = sX[(lane + 32 * h + 2) + UX_X * ((ly + 2) + UX_Y * (lz + 2))];
    }
  }
}
template <class OP>
__global__ void __launch_bounds__(256) k_mg_up0(LvlD F, LvlD C, float trivialScale, const float* __restrict__ xin, float* __restrict__ xout,
                                                int firstColor, int z0, int z1) {
  MF_PDL_PROLOGUE
  const int hx = (int)blockIdx.x * UP_X - 2, hy = (int)blockIdx.y * UP_Y - 2, hz = z0 + (int)blockIdx.z * UP_Z - 2;
  const bool interior = hx >= 0 && hy >= 0 && hz >= 0 && hx + UX_X <= F.sx && hy + UX_Y <= F.sy && hz + UX_Z <= F.sz;
  __shared__ float sX[UX_X * UX_Y * UX_Z];
  __shared__ float sC[UC_X * UC_Y * UC_Z];
  if (interior) up0Body<OP, false>(F, C, trivialScale, xin, xout, firstColor, z0, z1, sX, sC);
  else          up0Body<OP, true>(F, C, trivialScale, xin, xout, firstColor, z0, z1, sX, sC);
}

}  // namespace mf
