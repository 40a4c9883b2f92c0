// mg_wave0.cuh -- level-0 legs of the V(1,1)-cycle as z-marching wavefronts (3-D, 16-bit operator code, sx % 4 == 0).
//
// A CTA owns an xy tile and marches through a chunk of z planes.  Per step one new plane enters and the stages of the leg run
// on successive planes behind it, so every dependency is one plane old and sits in a 4-plane ring in shared memory:
//     down leg   plane p: first-colour values (x = 0 start: rhs/diag)      knSmoothColor (multigrid.cpp:668), colour 0
//                plane p-1: second colour from the six first-colour neighbours                        colour 1
//                plane p-2: residual r = b - A x                                                      knCalcResidual :739
//                plane p-3: its 9 weighted terms go into the coarse right-hand sides above/below      knRestrict :904
//     up leg     plane p: x + P x_coarse                                    knInterpolate :934 + knAddAssign
//                plane p-1: colour 1;  plane p-2: colour 0; store                                     knSmoothColor
// There is no halo in z inside a chunk and the xy halo (3 cells down, 2 up) is recomputed from b / the operator code / x, never
// read from a neighbour CTA's output.  The residual never reaches memory: 10.5 B/cell (b 4, code 2, x 4, coarse b 0.5) for the
// whole down leg instead of 28.5 for three passes, 14.5 for the up leg.
// A warp owns region rows, a lane the cell PAIR (2*lane, 2*lane+1) of a row: one cell of each colour, so no lane idles in a
// colour stage, and the pair's rhs and code are loaded once (8 + 4 bytes) and kept in registers for the three stages.
// Every value is computed by the same float expression as in mg_kernels.cuh (L0Code ops); cells outside the grid are inactive
// rows with x = +0, which leaves the guarded sums of the reference bit-unchanged.
#pragma once
#include "mg_kernels.cuh"
#include "multigrid.h"

namespace mf {

constexpr int WV_RX = 64, WV_RY = 32;          // region (tile + halo) of a CTA: 64 columns = 32 lane pairs, 32 rows = 4 per warp
constexpr int WV_NT = 512, WV_NW = WV_NT / 32;  // threads / warps per CTA
constexpr int WV_RPW = WV_RY / WV_NW;          // region rows per warp (row = warp + WV_NW * i)
constexpr int WV_HX = 4;                       // x halo (3 / 2 needed; 4 keeps the pairs 8-byte aligned)
constexpr int WV_TX = WV_RX - 2 * WV_HX;       // 56 owned columns
constexpr int WD_HY = 3, WD_TY = WV_RY - 2 * WD_HY;   // down leg: 26 owned rows
constexpr int WU_HY = 2, WU_TY = WV_RY - 2 * WU_HY;   // up leg: 28 owned rows

// Work distribution: a 1-D grid of exactly `waves x resident CTAs` items.  Every xy tile is cut into q or q+1 z chunks; the
// tiles with the longer chunks come first in launch order, so the shorter ones fill the tail.
struct WvSched { int tx, ty, q, nq; };          // tiles [0, nq) have q chunks, tiles [nq, tx*ty) have q+1
__device__ __forceinline__ void wvWhere(const WvSched& S, int U0, int U1, int& tileX, int& tileY, int& c0, int& c1) {
  const int id = blockIdx.x, first = S.nq * S.q;
  int tile, chunk, nch;
  if (id < first) { tile = id / S.q; chunk = id - tile * S.q; nch = S.q; }
  else { const int j = id - first; tile = j / (S.q + 1); chunk = j - tile * (S.q + 1); tile += S.nq; nch = S.q + 1; }
  tileX = tile % S.tx; tileY = tile / S.tx;
  const int planes = U1 - U0;                                      // chunk boundaries are even (U0 is)
  c0 = U0 + ((int)((long long)planes * chunk / nch) & ~1);
  c1 = (chunk + 1 == nch) ? U1 : U0 + ((int)((long long)planes * (chunk + 1) / nch) & ~1);
}

struct WvPair { float b0, b1; unsigned c; };    // rhs and operator codes of the two cells of a lane

__device__ __forceinline__ WvPair wvLoadPair(const LvlD& F, const float* __restrict__ b, int gx0, int gy, int gz, bool inXY) {
  WvPair r; r.b0 = 0.f; r.b1 = 0.f; r.c = 0u;
  if (inXY && gz >= 0 && gz < F.sz) {
    const int v = gx0 + F.sx * (gy + F.sy * gz);
    const float2 bb = *reinterpret_cast<const float2*>(b + v);
    r.b0 = bb.x; r.b1 = bb.y;
    r.c = *reinterpret_cast<const unsigned*>(F.code + v);
  }
  return r;
}
// the same from a precomputed in-plane offset (gx0 + sx * gy) plus the plane's offset: one 32-bit add per row and plane instead of
// the index product and four range tests (the loads were 10 % of the down leg's instructions)
__device__ __forceinline__ WvPair wvLoadPairAt(const float* __restrict__ b, const unsigned short* __restrict__ code, int v, bool ok) {
  WvPair r; r.b0 = 0.f; r.b1 = 0.f; r.c = 0u;
  if (ok) {
    const float2 bb = *reinterpret_cast<const float2*>(b + v);
    r.b0 = bb.x; r.b1 = bb.y;
    r.c = *reinterpret_cast<const unsigned*>(code + v);
  }
  return r;
}
__device__ __forceinline__ int wvCode(const WvPair& r, int w) { return w ? (int)(r.c >> 16) : (int)(r.c & 0xffffu); }
__device__ __forceinline__ float wvRhs(const WvPair& r, int w) { return w ? r.b1 : r.b0; }

// rhs - sum_offdiag(A x) in the reference's order (-x,+x,-y,+y,-z,+z; multigrid.cpp:686-689, :750-753)
__device__ __forceinline__ float wvOffdiag(const L0Code& op, const L0Code::Row& row, float b, float xm, float xp, float ym, float yp, float zm, float zp) {
  float sum = op.rhs(row, b);
  sum = op.template sub<0>(row, 0, 0, sum, xm);
  sum = op.template sub<1>(row, 0, 0, sum, xp);
  sum = op.template sub<2>(row, 0, 0, sum, ym);
  sum = op.template sub<3>(row, 0, 0, sum, yp);
  sum = op.template sub<4>(row, 0, 0, sum, zm);
  sum = op.template sub<5>(row, 0, 0, sum, zp);
  return sum;
}
// the lane's pair (columns 2*lane, 2*lane+1) of one ring row and of its four y/z neighbour rows: conflict-free 8-byte loads
struct WvNbr { float2 c, ym, yp, zm, zp; float left, right; };   // left = column 2*lane-1, right = column 2*lane+2 (by shuffle)
__device__ __forceinline__ WvNbr wvNeighbours(const float (*xs)[WV_RY][WV_RX], int q, int ry, int lane) {
  WvNbr n;
  n.c  = *reinterpret_cast<const float2*>(&xs[q & 3][ry][2 * lane]);
  n.ym = *reinterpret_cast<const float2*>(&xs[q & 3][ry - 1][2 * lane]);
  n.yp = *reinterpret_cast<const float2*>(&xs[q & 3][ry + 1][2 * lane]);
  n.zm = *reinterpret_cast<const float2*>(&xs[(q - 1) & 3][ry][2 * lane]);
  n.zp = *reinterpret_cast<const float2*>(&xs[(q + 1) & 3][ry][2 * lane]);
  n.left = __shfl_up_sync(0xffffffffu, n.c.y, 1);
  n.right = __shfl_down_sync(0xffffffffu, n.c.x, 1);
  return n;
}
// the same with the six neighbours read from the ring: xc points at the cell in its plane, xm / xp at the same cell one plane
// below / above (used where a lane relaxes ONE cell of its pair)
__device__ __forceinline__ float wvOffdiagAt(const L0Code& op, const L0Code::Row& row, float b, const float* xc, const float* xm, const float* xp) {
  return wvOffdiag(op, row, b, xc[-1], xc[1], xc[-WV_RX], xc[WV_RX], *xm, *xp);
}

// ---- down leg ---------------------------------------------------------------------------------------------------------------
// Chunk [c0, c1) of fine planes (c0 even): the CTA writes x on the planes of the chunk inside the rank's [z0, z1), and the
// coarse right-hand side (and x_coarse = 0) of the coarse vertices (X, Y, K) with 2X, 2Y in its tile, 2K in the chunk and K in
// the rank's coarse range [cz0, cz1).  Tiles cover [0, sx] x [0, sy] and the chunks [U0, U1) reach plane sz when the last coarse
// plane needs it, so the coarse vertices at 2X == sx (even sizes) have an owner too.
__global__ void __launch_bounds__(WV_NT, 2) k_mg_down0(LvlD F, LvlD C, float trivialScale, int firstColor, int U0, int U1, WvSched S,
                                                  int z0, int z1, int cz0, int cz1, int foldFirst, const int* __restrict__ skip) {
  pdlTrigger();              // griddepcontrol.wait follows the set-up (ring clears, tile geometry): it overlaps the predecessor's drain
  __shared__ __align__(16) float xs[4][WV_RY][WV_RX];     // read and written as float2
  __shared__ __align__(16) float rs[WV_RY][WV_RX];
  const L0Code op(F, trivialScale);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int tileX, tileY, c0, c1;
  wvWhere(S, U0, U1, tileX, tileY, c0, c1);
  const int ox = tileX * WV_TX, oy = tileY * WD_TY;
  const int hx = ox - WV_HX, hy = oy - WD_HY;
  const int gx0 = hx + 2 * lane;                                  // even: (gx0 + gy + gz) & 1 == (gy + gz) & 1
  const bool inX = gx0 >= 0 && gx0 < F.sx;
  for (int t = threadIdx.x; t < WV_RY * WV_RX; t += WV_NT) reinterpret_cast<float4*>(&xs[0][0][0])[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int t = threadIdx.x; t < WV_RY * WV_RX / 4; t += WV_NT) reinterpret_cast<float4*>(&rs[0][0])[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  // coarse vertices of the tile
  const int CX0 = ox >> 1, CY0 = oy >> 1;
  const int ncx = min(WV_TX / 2, C.sx - CX0), ncy = min(WD_TY / 2, C.sy - CY0);
  const int nslots = (ncx > 0 && ncy > 0) ? ncx * ncy : 0;
  constexpr int NSLOT = (WV_TX / 2 * (WD_TY / 2) + WV_NT - 1) / WV_NT;   // coarse vertices per thread
  float acc[NSLOT] = {};
  float sumB[WV_RPW] = {}, sumC[WV_RPW] = {};   // rhs - offdiag of the second-colour cell (planes p-1, p-2)
  WvPair rowA[WV_RPW], rowB[WV_RPW], rowC[WV_RPW], rowN[WV_RPW];                      // planes p, p-1, p-2 and the prefetch of p+1
  int rowOff[WV_RPW]; unsigned rowOk = 0u;                            // in-plane offset of the lane's pair in each of its rows; rows inside the grid
  const int planeCells = F.sx * F.sy;
#pragma unroll
  for (int i = 0; i < WV_RPW; i++) {
    rowB[i].b0 = rowB[i].b1 = 0.f; rowB[i].c = 0u; rowC[i] = rowB[i];
    const int gy = hy + warp + WV_NW * i;
    rowOff[i] = gx0 + F.sx * gy;
    if (inX && gy >= 0 && gy < F.sy) rowOk |= 1u << i;
  }
  pdlWait();
  // launched ahead of the caller's decision (mg_apply_begin): the PCG has converged, this V-cycle is not wanted
  if (skip && *reinterpret_cast<const volatile int*>(skip)) return;
  {
    const int gz = c0 - 3; const bool zok = gz >= 0 && gz < F.sz;
#pragma unroll
    for (int i = 0; i < WV_RPW; i++) rowA[i] = wvLoadPairAt(F.b, F.code, rowOff[i] + planeCells * gz, zok && ((rowOk >> i) & 1u));
  }
  __syncthreads();
  const int pEnd = c1 + 2 + (c1 & 1);
  for (int p = c0 - 3; p <= pEnd; p++) {
    {                                                                   // prefetch: consumed at the end of the step
      const int gz = p + 1; const bool zok = gz >= 0 && gz < F.sz && p < pEnd; const int po = planeCells * gz;
#pragma unroll
      for (int i = 0; i < WV_RPW; i++) rowN[i] = wvLoadPairAt(F.b, F.code, rowOff[i] + po, zok && ((rowOk >> i) & 1u));
    }
    // ---- plane p: first-colour values (x = 0: no neighbour terms)
#pragma unroll
    for (int i = 0; i < WV_RPW; i++) {
      const int ry = warp + WV_NW * i, gy = hy + ry;
      const int w = (gy + p + firstColor) & 1;                    // which cell of the pair has the first colour
      L0Code::Row r; r.c = wvCode(rowA[i], w);
      float v = 0.f;
      if (!op.inactive(r)) v = op.rhs(r, wvRhs(rowA[i], w)) / op.diag(r, 0);
      xs[p & 3][ry][2 * lane + w] = v;
    }
    // ---- plane q = p-3: restriction terms of its residual (in rs since the previous step)
    {
      const int q = p - 3;
      const float wz = (q & 1) ? 0.5f : 1.0f;
#pragma unroll
      for (int jj = 0; jj < NSLOT; jj++) {
        const int slot = threadIdx.x + WV_NT * jj;
        if (slot < nslots) {
          const int X = slot % ncx, Y = slot / ncx;
          const float* rc = &rs[2 * (CY0 + Y) - hy][2 * (CX0 + X) - hx];
          float t[9];
#pragma unroll
          for (int dy = -1; dy <= 1; dy++)
#pragma unroll
            for (int dx = -1; dx <= 1; dx++)
              t[(dy + 1) * 3 + dx + 1] = (wz * ((dx ? 0.5f : 1.0f) * (dy ? 0.5f : 1.0f))) * rc[dy * WV_RX + dx];   // rw = 1/2^(odd coordinates), exact
          float a = acc[jj];
#pragma unroll
          for (int e = 0; e < 9; e++) a += t[e];
          if (q & 1) {
            // q = 2K+1 closes coarse plane K and opens K+1 (q = 2(K+1)-1)
            const int K = (q - 1) >> 1;
            if (2 * K >= c0 && 2 * K < c1 && K >= cz0 && K < cz1) {
              const int v = mgIdx(C, CX0 + X, CY0 + Y, K);
              float x0 = 0.f;
              if (!mgRowZero(C, v)) {                                  // inactive coarse vertex (all-zero row): b stays 0
                C.b[v] = a;
                x0 = mgFoldFirst(C, v, CX0 + X, CY0 + Y, K, a, foldFirst != 0);
              }
              C.x[v] = x0;
            }
            a = 0.f;
#pragma unroll
            for (int e = 0; e < 9; e++) a += t[e];
          }
          acc[jj] = a;
        }
      }
    }
    __syncthreads();
    // ---- plane q = p-1: second colour
    {
      const int q = p - 1;
#pragma unroll
      for (int i = 0; i < WV_RPW; i++) {
        const int ry = warp + WV_NW * i, gy = hy + ry;
        if (ry < 1 || ry > WV_RY - 2) continue;                    // warp-uniform
        const int w = 1 - ((gy + q + firstColor) & 1);
        const int col = 2 * lane + w;
        if (col >= 1 && col <= WV_RX - 2) {
          L0Code::Row r; r.c = wvCode(rowB[i], w);
          float v = 0.f;
          if (!op.inactive(r)) {
            sumB[i] = wvOffdiagAt(op, r, wvRhs(rowB[i], w), &xs[q & 3][ry][col], &xs[(q - 1) & 3][ry][col], &xs[(q + 1) & 3][ry][col]);
            v = sumB[i] / op.diag(r, 0);
          }
          xs[q & 3][ry][col] = v;
        }
      }
    }
    __syncthreads();
    // ---- plane q = p-2: residual (both colours are final on planes q-1, q, q+1), store x
    {
      const int q = p - 2;
      const bool storeQ = q >= c0 && q < c1 && q >= z0 && q < z1;
#pragma unroll
      for (int i = 0; i < WV_RPW; i++) {
        const int ry = warp + WV_NW * i, gy = hy + ry;
        if (ry < 2 || ry > WV_RY - 3) continue;                    // warp-uniform
        const WvNbr n = wvNeighbours(xs, q, ry, lane);
        if (lane < 1 || lane > 30) continue;
        const int w2 = 1 - ((gy + q + firstColor) & 1);             // the second-colour cell of the pair
        float res[2];
#pragma unroll
        for (int w = 0; w < 2; w++) {
          L0Code::Row r; r.c = wvCode(rowC[i], w);
          float sum = 0.f;                                          // inactive vertices: r = 0
          if (!op.inactive(r)) {
            // the second-colour cell was relaxed from exactly these (final) neighbour values: its rhs - offdiag sum is sumC
            if (w == w2) sum = sumC[i];
            else sum = w ? wvOffdiag(op, r, rowC[i].b1, n.c.x, n.right, n.ym.y, n.yp.y, n.zm.y, n.zp.y)
                         : wvOffdiag(op, r, rowC[i].b0, n.left, n.c.y, n.ym.x, n.yp.x, n.zm.x, n.zp.x);
            sum -= op.diag(r, 0) * (w ? n.c.y : n.c.x);
          }
          res[w] = sum;
        }
        *reinterpret_cast<float2*>(&rs[ry][2 * lane]) = make_float2(res[0], res[1]);
        if (storeQ && ry >= WD_HY && ry < WD_HY + WD_TY && lane >= WV_HX / 2 && lane < (WV_HX + WV_TX) / 2 && inX && gy < F.sy)
          *reinterpret_cast<float2*>(F.x + gx0 + F.sx * (gy + F.sy * q)) = n.c;
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < WV_RPW; i++) { rowC[i] = rowB[i]; rowB[i] = rowA[i]; rowA[i] = rowN[i]; sumC[i] = sumB[i]; }
  }
}

// ---- up leg -----------------------------------------------------------------------------------------------------------------
// xout = smooth(colour 1 - firstColor... second)(smooth(firstColor)(xin + P x_coarse)) on the planes of the chunk inside [z0, z1).
// xin is the down leg's x (another array than xout: the halo planes of a neighbour chunk must keep their old values).
constexpr int WU_CX = WV_RX / 2 + 1, WU_CY = WV_RY / 2 + 1;        // coarse vertices under the region: 33 x 17

// sigmaNew = xout . b, beta = sigmaNew / sigma (FinBeta of cg.cuh on the caller's scalars)
struct WvFinBeta {
  MgDotBeta d;
  __device__ void operator()(double* r) const {
    if (*d.done) return;
    const float sigmaNew = (float)r[0];
    *d.beta = sigmaNew / *d.sigma;
    *d.sigma = sigmaNew;
    if (!(*d.resNorm < 1e35f)) { *d.diverged = 1; *d.done = 1; }
  }
};
// DOT: the kernel also reduces sum (double)(xout * b) over the cells it stores -- the CG's next dot product (GridDotProduct,
// conjugategrad.cpp:175: float products accumulated in double) -- while both factors are in registers
template <bool DOT>
__global__ void __launch_bounds__(WV_NT, 2) k_mg_up0w(LvlD F, LvlD C, float trivialScale, int firstColor, const float* __restrict__ xin,
                                                 float* __restrict__ xout, int U0, int U1, WvSched S, int z0, int z1, WvFinBeta fin, RedCtx ctx) {
  pdlTrigger();              // griddepcontrol.wait follows the set-up, in front of the first load of data a predecessor wrote
  double dotAcc[1] = { 0.0 };
  __shared__ __align__(16) float xs[4][WV_RY][WV_RX];     // read and written as float2
  __shared__ __align__(16) float cs[2][WU_CY][WU_CX];
  const L0Code op(F, trivialScale);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int tileX, tileY, c0, c1;
  wvWhere(S, U0, U1, tileX, tileY, c0, c1);
  const int ox = tileX * WV_TX, oy = tileY * WU_TY;
  const int hx = ox - WV_HX, hy = oy - WU_HY;
  const int gx0 = hx + 2 * lane;
  const bool inX = gx0 >= 0 && gx0 < F.sx;
  const int CX0 = hx >> 1, CY0 = hy >> 1;                          // arithmetic shift: floor (hx, hy are even)
  for (int t = threadIdx.x; t < WV_RY * WV_RX; t += WV_NT) reinterpret_cast<float4*>(&xs[0][0][0])[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  // coarse plane K -> cs[K & 1]; x_coarse is 0 at inactive vertices and in the padding, nothing exists beyond the padding
  auto loadCoarse = [&](int K) {
    for (int t = threadIdx.x; t < WU_CY * WU_CX; t += WV_NT) {
      const int X = CX0 + t % WU_CX, Y = CY0 + t / WU_CX;
      cs[K & 1][t / WU_CX][t % WU_CX] = (X >= -1 && Y >= -1 && K >= -1 && X <= C.sx && Y <= C.sy && K <= C.sz) ? C.x[mgIdx(C, X, Y, K)] : 0.f;
    }
  };
  const int pBeg = c0 - 2, pEnd = c1 + 1;
  pdlWait();
  loadCoarse(pBeg >> 1);
  if (pBeg & 1) loadCoarse((pBeg + 1) >> 1);
  WvPair rowA[WV_RPW], rowB[WV_RPW], rowC[WV_RPW], rowN[WV_RPW];
  float2 xinA[WV_RPW], xinN[WV_RPW];
  int rowOff[WV_RPW]; unsigned rowOk = 0u;                            // in-plane offset of the lane's pair in each of its rows; rows inside the grid
  const int planeCells = F.sx * F.sy;
  auto loadXAt = [&](int v, bool ok) {
    float2 r = make_float2(0.f, 0.f);
    if (ok) r = *reinterpret_cast<const float2*>(xin + v);
    return r;
  };
#pragma unroll
  for (int i = 0; i < WV_RPW; i++) {
    rowB[i].b0 = rowB[i].b1 = 0.f; rowB[i].c = 0u; rowC[i] = rowB[i];
    const int gy = hy + warp + WV_NW * i;
    rowOff[i] = gx0 + F.sx * gy;
    if (inX && gy >= 0 && gy < F.sy) rowOk |= 1u << i;
  }
  {
    const bool zok = pBeg >= 0 && pBeg < F.sz;
#pragma unroll
    for (int i = 0; i < WV_RPW; i++) {
      const bool ok = zok && ((rowOk >> i) & 1u);
      rowA[i] = wvLoadPairAt(F.b, F.code, rowOff[i] + planeCells * pBeg, ok);
      xinA[i] = loadXAt(rowOff[i] + planeCells * pBeg, ok);
    }
  }
  __syncthreads();
  for (int p = pBeg; p <= pEnd; p++) {
    {                                                                   // prefetch of plane p+1
      const int gz = p + 1; const bool zok = gz >= 0 && gz < F.sz && p < pEnd; const int po = planeCells * gz;
#pragma unroll
      for (int i = 0; i < WV_RPW; i++) {
        const bool ok = zok && ((rowOk >> i) & 1u);
        rowN[i] = wvLoadPairAt(F.b, F.code, rowOff[i] + po, ok);
        xinN[i] = loadXAt(rowOff[i] + po, ok);
      }
    }
    // ---- plane p: x + P x_coarse (knInterpolate's loop nest z, y, x over [v/2, (v+1)/2]; absent terms add +0)
    {
      const int pz = p & 1;
      const float (*cz0)[WU_CX] = cs[(p >> 1) & 1];
      const float (*cz1)[WU_CX] = cs[((p + 1) >> 1) & 1];
#pragma unroll
      for (int i = 0; i < WV_RPW; i++) {
        const int ry = warp + WV_NW * i, gy = hy + ry;
        const int py = gy & 1;
        const int cy = (gy >> 1) - CY0, cx = lane;                 // (gx0 >> 1) - CX0 == lane
        float out[2];
#pragma unroll
        for (int w = 0; w < 2; w++) {                              // w == x parity of the cell
          float val = w ? xinA[i].y : xinA[i].x;
          L0Code::Row r; r.c = wvCode(rowA[i], w);
          if (!op.inactive(r)) {
            float sum = 0.f;
            sum += cz0[cy][cx];
            sum += w ? cz0[cy][cx + 1] : 0.f;
            sum += py ? cz0[cy + 1][cx] : 0.f;
            sum += (w & py) ? cz0[cy + 1][cx + 1] : 0.f;
            sum += pz ? cz1[cy][cx] : 0.f;
            sum += (pz & w) ? cz1[cy][cx + 1] : 0.f;
            sum += (pz & py) ? cz1[cy + 1][cx] : 0.f;
            sum += (pz & py & w) ? cz1[cy + 1][cx + 1] : 0.f;
            val += __int_as_float((127 - (w + py + pz)) << 23) * sum;      // iw = 1 / 2^(odd coordinates), exact
          }
          out[w] = val;
        }
        *reinterpret_cast<float2*>(&xs[p & 3][ry][2 * lane]) = make_float2(out[0], out[1]);
      }
    }
    __syncthreads();
    // ---- plane q = p-1: first colour of the post sweep
    {
      const int q = p - 1;
#pragma unroll
      for (int i = 0; i < WV_RPW; i++) {
        const int ry = warp + WV_NW * i, gy = hy + ry;
        if (ry < 1 || ry > WV_RY - 2) continue;                    // warp-uniform
        const int w = (gy + q + firstColor) & 1;
        const int col = 2 * lane + w;
        if (col >= 1 && col <= WV_RX - 2) {
          L0Code::Row r; r.c = wvCode(rowB[i], w);
          if (!op.inactive(r))
            xs[q & 3][ry][col] = wvOffdiagAt(op, r, wvRhs(rowB[i], w), &xs[q & 3][ry][col], &xs[(q - 1) & 3][ry][col], &xs[(q + 1) & 3][ry][col]) / op.diag(r, 0);
        }
      }
    }
    __syncthreads();
    // ---- plane q = p-2: second colour, store; fetch the coarse plane the next step needs
    {
      const int q = p - 2;
      const bool storeQ = q >= c0 && q < c1 && q >= z0 && q < z1;
#pragma unroll
      for (int i = 0; i < WV_RPW; i++) {
        const int ry = warp + WV_NW * i, gy = hy + ry;
        if (ry < WU_HY || ry >= WU_HY + WU_TY) continue;           // warp-uniform
        const int w = 1 - ((gy + q + firstColor) & 1);
        if (lane < WV_HX / 2 || lane >= (WV_HX + WV_TX) / 2) continue;
        const int col = 2 * lane + w;
        L0Code::Row r; r.c = wvCode(rowC[i], w);
        float2 out = *reinterpret_cast<const float2*>(&xs[q & 3][ry][2 * lane]);
        if (!op.inactive(r)) {
          const float v = wvOffdiagAt(op, r, wvRhs(rowC[i], w), &xs[q & 3][ry][col], &xs[(q - 1) & 3][ry][col], &xs[(q + 1) & 3][ry][col]) / op.diag(r, 0);
          if (w) out.y = v; else out.x = v;
        }
        if (storeQ && inX && gy < F.sy) {
          *reinterpret_cast<float2*>(xout + gx0 + F.sx * (gy + F.sy * q)) = out;
          if (DOT) { dotAcc[0] += (double)(out.x * rowC[i].b0); dotAcc[0] += (double)(out.y * rowC[i].b1); }
        }
      }
      if (p < pEnd && ((p + 1) & 1)) loadCoarse((p + 2) >> 1);     // plane p+1 odd: it also needs coarse plane (p+2)/2
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < WV_RPW; i++) { rowC[i] = rowB[i]; rowB[i] = rowA[i]; rowA[i] = rowN[i]; xinA[i] = xinN[i]; }
  }
  if (DOT) {
    __shared__ double sh[32];
    const int ops[1] = { 0 };
    gridReduceK<1>(dotAcc, ops, ctx, sh, fin);
  }
}

}  // namespace mf
