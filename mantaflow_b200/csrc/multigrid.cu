// multigrid.cu -- GridMg on the device: hierarchy construction (setA) and V-cycle (doVCycle).
// Reference: source/multigrid.cpp (ctor :220-319, setA :386-415, doVCycle :448-504, genCoarseGrid :520-578,
// knGenCoarseGridOperator :580-657, smoothGS :713-737, solveCG :796-902).
#include "multigrid.h"
#include "mg_kernels.cuh"
#include "mg_fused0.cuh"
#include "mg_wave0.cuh"
#include "mg_tail.cuh"
#include "comm.h"
#include <algorithm>
#include <cmath>
#include <cstring>

namespace mf {

struct MgHierarchy {
  Solver* s;
  MgDims D;
  std::vector<LvlD> lv;
  unsigned short* code0;    // level 0: compact operator (allocated on first use)
  bool useCode;             // level-0 kernels read code0 instead of float planes
  bool aSet;
  bool hasTrivial, nonZeroStencilSum;
  float coarsestAcc, trivialScale;
  int pre, post;
  MgPath* dPaths; int npaths;
  int* dFlags;              // job-wide device scratch: [0..3] activate flags, [4..5] coarsening flags
  int usedFallback;
  bool pendingSolo;         // a group of rank-0-only launches has not been followed by a cross-rank barrier yet
  std::vector<std::vector<int>> pairSplit;   // blocked level l: rank r sweeps z pairs [pairSplit[l][r], pairSplit[l][r+1])
  std::vector<MgOff8> offs;                  // blocked level l: stencil offset tables of the sweeps (buildOffsets)
  __half* half1; long long ash1;             // level 1 (3-D): storage for the binary16 operator (LvlD::Ah points here when used)
  int nspecial1;                             // level-1 vertices that keep the float planes (k_mg_mark_special)
  int* special1;                             // job-wide: count[8], list[8][MG_SPECIAL_CAP] of those vertices by colour
  int* special1Local;                        // this process' copy (the sweeps read it in every launch)
  // persistent kernels for the small levels (mg_tail.cuh): grid-barrier counter, its host-tracked value, time-out flag
  int tailCluster; unsigned* tailBar; unsigned tailBase; int* tailErrHost; int* tailErrDev; MgOff8* tailOffs;
  int tailFirst;            // first level swept by the tail kernels, or -1
  // the last swept level + the coarsest level in one CTA (mg_bottom.cuh)
  bool bottomOn; MgOff8* bottomOff; size_t bottomSmem; int bottomVec;
  // mg_apply_begin: the level-0 down leg of the next mg_apply is already queued
  bool begun; float* begunDst; const float* begunRhs; float* begunR0;
};

// ---- partition of a level over the ranks ----------------------------------------------------------------------
// Levels are split by z (level 0: the solver's slabs; blocked levels: pairs of z planes).  Levels too small to be worth
// splitting are handled by rank 0 alone ("solo"), and consecutive solo launch groups need no cross-rank barrier in between.
static bool lvlSolo(const MgHierarchy* mg, int l) {
  const Solver* s = mg->s;
  if (!s->multi() || l == 0) return false;
  const LvlD& L = mg->lv[l];
  // below ~1M vertices a colour sweep takes about as long as one cross-GPU barrier
  return L.n < (1 << 20) || (L.PZ - 2) < 2 * s->nranks;
}
// blocked level: range of z pairs of rank r
static void lvlPairs(const MgHierarchy* mg, int l, int r, int& lz0, int& lz1) {
  const Solver* s = mg->s; const LvlD& L = mg->lv[l];
  const int BZ = mg->D.is3D ? L.PZ - 2 : 1;
  if (!s->multi()) { lz0 = 0; lz1 = BZ; return; }
  lz0 = mg->pairSplit[l][r]; lz1 = mg->pairSplit[l][r + 1];
}
// natural z range of rank r on level l
static void lvlPlanes(const MgHierarchy* mg, int l, int r, int& z0, int& z1) {
  const Solver* s = mg->s; const LvlD& L = mg->lv[l];
  if (l == 0) { z0 = s->multi() ? s->zsplit[r] : 0; z1 = s->multi() ? s->zsplit[r + 1] : L.sz; return; }
  int a, b; lvlPairs(mg, l, r, a, b);
  z0 = 2 * a < L.sz ? 2 * a : L.sz; z1 = 2 * b < L.sz ? 2 * b : L.sz;
}
// storage range of rank r on level l (setup kernels)
static void lvlStore(const MgHierarchy* mg, int l, int r, int& p0, int& p1) {
  const Solver* s = mg->s; const LvlD& L = mg->lv[l];
  if (!s->multi()) { p0 = 0; p1 = L.na; return; }
  if (lvlSolo(mg, l)) { p0 = 0; p1 = (r == 0) ? L.na : 0; return; }
  if (l == 0) { const long long xy = (long long)L.sx * L.sy; p0 = (int)(xy * s->zsplit[r]); p1 = (int)(xy * s->zsplit[r + 1]); return; }
  const int pair = L.NC * L.PY * L.PX;            // storage entries per z pair
  p0 = (r == 0) ? 0 : (mg->pairSplit[l][r] + L.zoff) * pair;
  p1 = (r == s->nranks - 1) ? L.na : (mg->pairSplit[l][r + 1] + L.zoff) * pair;
}
static void beginGroup(MgHierarchy* mg, bool solo) { if (!solo && mg->pendingSolo) { mg->s->rankSync(); mg->pendingSolo = false; } }
static void endGroup(MgHierarchy* mg, bool solo) { if (!solo) mg->s->rankSync(); else if (mg->s->multi()) mg->pendingSolo = true; }
static void flushGroups(MgHierarchy* mg) { if (mg->pendingSolo) { mg->s->rankSync(); mg->pendingSolo = false; } }

// ---- coarsest level: Jacobi-preconditioned CG in double, one CTA (solveCG, multigrid.cpp:796-902) ---------
// One thread per vertex (n <= 1024).  The 27 (7) coefficients of the thread's row are loaded once into registers, the
// search vector lives in shared memory; 3 block barriers per iteration.  An iteration is a chain of dependent double
// instructions on one SM (measured per phase with clock64, 736 threads, 9^3 level of a 256^3 grid, cycles: row 2100 -> 1500,
// p.Ap sum 1060 -> 870, update 340 -> 120, r.r / r.z sums 1850 -> 1490):
//  * the row: 27 LDS.64 of the search vector per thread are 157 KB per iteration = 1230 cycles of the SM's shared-memory
//    port, the floor of this phase.  The coefficients stay 27 floats in registers and are converted where they are used: as
//    54 registers of doubles ptxas has no room left to batch the loads and serialises LDS -> DFMA -> LDS on one register
//    pair.  No mask test per term: a term that does not exist has coefficient 0 and reads a zeroed margin of the vector;
//  * a term is one DFMA (explicit __fma_rn in this -fmad=false translation unit; the reference's own sums are sequential
//    doubles whose order the block-wide trees below do not keep either -- the level is solved to 1e-4 of the outer accuracy,
//    the difference is ~1e-16 relative), the Jacobi step multiplies by a reciprocal computed once;
//  * block-wide sums of one or two values: warp tree, one partial per warp in shared memory, then warp 0 alone combines the
//    partials and runs `fin` -- the scalar double divisions / square root of the iteration (130 / 105 cycles each), which
//    would otherwise occupy the FP64 pipe once per warp -- before the results are published through shared memory.  Two
//    block barriers per call.
template <int NV, class Fin>
__device__ __forceinline__ void cgBlockSum(double a, double b, double (*sh)[32], double* out, Fin fin) {
  for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); if (NV > 1) b += __shfl_xor_sync(0xffffffffu, b, o); }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sh[0][w] = a; if (NV > 1) sh[1][w] = b; }
  __syncthreads();
  if (w == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    a = (l < nw) ? sh[0][l] : 0.0;
    if (NV > 1) b = (l < nw) ? sh[1][l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); if (NV > 1) b += __shfl_xor_sync(0xffffffffu, b, o); }
    if (l == 0) fin(a, b, out);
  }
  __syncthreads();
}

template <bool LEVEL0>
__device__ __forceinline__ void mgCoarseCg(const LvlD& L, const MgDims& D, float accuracy, double (*sh)[32], double* out, double* vecBase,
                                           bool fromReg = false, float bReg = 0.f, float* xShared = nullptr) {
  // fromReg (mg_bottom.cuh): the right-hand side arrives in a register, the initial guess is 0, and the solution also goes to
  // xShared (the level's x in shared memory) -- what the restriction kernel would have left in L.b / L.x
  constexpr int NS = LEVEL0 ? 7 : 27;
  const int v = threadIdx.x;
  const bool in = v < L.n;
  const int vx = in ? v % L.sx : 0, vy = in ? (v % (L.sx * L.sy)) / L.sx : 0, vz = in ? v / (L.sx * L.sy) : 0;
  const int p = in ? mgIdx(L, vx, vy, vz) : 0;
  const bool active = in && L.type[p] != vtInactive;
  // row of A in the order the reference applies it; a term that does not exist (neighbour outside the grid or inactive, or an
  // inactive row) has coefficient 0 and reads a finite value: vec has a zeroed margin of Z + Y + 1 entries on both sides
  float coef[NS];
#pragma unroll
  for (int s = 0; s < NS; s++) coef[s] = 0.f;
  const int Y = L.sx, Z = L.sx * L.sy;
  const int M = Z + Y + 1;
  double* vec = vecBase + M;
  for (int t = threadIdx.x; t < M; t += blockDim.x) { vecBase[t] = 0.0; vec[blockDim.x + t] = 0.0; }
  if (active) {
    if (LEVEL0) {
      // order -x,+x,-y,+y,-z,+z, centre (multigrid.cpp:840-850 for l == 0)
      const long long n = L.as;
      if (vx > 0)        coef[0] = L.A[n + v - 1];
      if (vx < L.sx - 1) coef[1] = L.A[n + v];
      if (vy > 0)        coef[2] = L.A[2 * n + v - Y];
      if (vy < L.sy - 1) coef[3] = L.A[2 * n + v];
      if (D.is3D) {
        if (vz > 0)        coef[4] = L.A[3 * n + v - Z];
        if (vz < L.sz - 1) coef[5] = L.A[3 * n + v];
      }
      coef[6] = L.A[v];
    } else {
      const int SS = D.SS;
      int s = 0;
#pragma unroll
      for (int dz = -1; dz <= 1; dz++)
#pragma unroll
        for (int dy = -1; dy <= 1; dy++)
#pragma unroll
          for (int dx = -1; dx <= 1; dx++) {
            if (dz < D.zmin || dz > D.zmax) continue;
            const int Nx = vx + dx, Ny = vy + dy, Nz = vz + dz;
            const int slot = (dx + 1) + 3 * (dy + 1) + 9 * (dz + 1);     // register slot (compile time)
            if (mgInGrid(L, Nx, Ny, Nz)) {
              const int pn = mgIdx(L, Nx, Ny, Nz);
              if (L.type[pn] != vtInactive) {
                coef[slot] = (s < SS) ? L.A[(SS - 1 - s) * L.as + pn] : L.A[(s - SS + 1) * L.as + p];
              }
            }
            s++;
          }
    }
  }
  // the loads of a row are issued in batches of 9 (one z plane of the stencil), the sum keeps the reference's term order
  auto applyA = [&]() {
    double sum = 0;
    constexpr int NB = LEVEL0 ? 7 : 9;
#pragma unroll
    for (int g = 0; g < NS; g += NB) {
      double xv[NB];
#pragma unroll
      for (int i = 0; i < NB; i++) {
        const int s = g + i;
        int off;
        if (LEVEL0) off = (s == 0) ? -1 : (s == 1) ? 1 : (s == 2) ? -Y : (s == 3) ? Y : (s == 4) ? -Z : (s == 5) ? Z : 0;
        else off = (s % 3 - 1) + Y * ((s / 3) % 3 - 1) + Z * (s / 9 - 1);
        xv[i] = vec[v + off];
      }
#pragma unroll
      for (int i = 0; i < NB; i++) {
        float c = coef[g + i];
        asm volatile("" : "+f"(c));          // keep the row as 27 floats: converted copies would not fit the register budget
        sum = __fma_rn((double)c, xv[i], sum);
      }
    }
    return sum;
  };
  const double invDiag = active ? 1.0 / (double)L.A[p] : 0.0;
  // everything above reads the operator only (constant between hierarchy builds): the wait for the predecessor comes here
  pdlWait();
  double x = (in && !fromReg) ? (double)L.x[p] : 0.0, r = 0.0, z = 0.0, pp = 0.0;
  vec[v] = x;
  __syncthreads();
  if (active) {
    r = (double)(fromReg ? bReg : L.b[p]) - applyA();
    z = r * invDiag;
    pp = z;
  }
  // out[0] = sqrt(sum r.r), out[1] = sum r.z   (inactive rows: r = z = pp = 0 throughout, their coefficients are all zero)
  cgBlockSum<2>(r * r, r * z, sh, out, [](double rr, double rz, double* o) { o[0] = sqrt(rr); o[1] = rz; });
  const double initialResidual = out[0];
  double alphaTop = out[1];
  const double acc = (double)accuracy;
#ifdef MF_CG_DBG
  long long tc[6] = {0, 0, 0, 0, 0, 0}; int nit = 0; const long long tstart = clock64();
#define TCK(i) { const long long _t = clock64(); tc[i] += _t - tl; tl = _t; }
#else
#define TCK(i)
#endif
  for (int iter = 0; iter < 10000 && initialResidual > 1E-12; iter++) {
#ifdef MF_CG_DBG
    long long tl = clock64(); nit++;
#endif
    vec[v] = pp;              // the previous contents were last read before the barriers of the reduction above
    __syncthreads();
    TCK(0)
    z = applyA();
    TCK(1)
    cgBlockSum<1>(pp * z, 0.0, sh, out, [alphaTop](double pz, double, double* o) { o[0] = alphaTop / pz; });
    TCK(2)
    const double alpha = out[0];
    x = __fma_rn(alpha, pp, x); r = __fma_rn(-alpha, z, r); z = r * invDiag;
    TCK(3)
    // out[0] = beta, out[1] = new alphaTop, out[2] = converged
    cgBlockSum<2>(r * r, r * z, sh, out, [alphaTop, initialResidual, acc](double rr, double rz, double* o) {
      o[0] = rz / alphaTop; o[1] = rz; o[2] = (sqrt(rr) / initialResidual < acc) ? 1.0 : 0.0; });
    TCK(4)
    if (out[2] != 0.0) break;
    const double beta = out[0];
    alphaTop = out[1];
    pp = __fma_rn(beta, pp, z);
  }
#ifdef MF_CG_DBG
  if (threadIdx.x == 0) printf("cgdbg it %d publish %lld row %lld sum1 %lld upd %lld sum2 %lld total %lld\n", nit, tc[0] / nit, tc[1] / nit, tc[2] / nit, tc[3] / nit, tc[4] / nit, clock64() - tstart);
#endif
  if (in) { L.x[p] = (float)x; if (xShared) xShared[p] = (float)x; }
}
} // namespace mf
#include "mg_bottom.cuh"
namespace mf {
// MAXT = 768: up to 85 registers per thread; with 1024 threads (64 registers) part of the state spills to local memory
template <bool LEVEL0, int MAXT>
__global__ void __launch_bounds__(MAXT) k_mg_coarse_cg(LvlD L, MgDims D, float accuracy) {
  pdlTrigger();                 // (griddepcontrol.wait inside mgCoarseCg, behind the loads of the row)
  __shared__ double sh[2][32];
  __shared__ double out[4];
  __shared__ double vec[1024 + 2 * (1024 + 1024 + 1)];      // n + two margins of Z + Y + 1 <= 2049
  mgCoarseCg<LEVEL0>(L, D, accuracy, sh, out, vec);
}

// ---- host side ----------------------------------------------------------------------------------------------
static void planTail(MgHierarchy* mg);
static void planBottom(MgHierarchy* mg);
static void buildPaths(MgHierarchy* mg) {
  // restatement of the path enumeration in GridMg::GridMg (multigrid.cpp:286-318)
  struct P { int N[3], U[3], W[3]; int sc, sf; float rw, iw; bool inU; };
  std::vector<P> paths;
  const int dim = mg->D.dim;
  const int smin[3] = { -1, -1, mg->D.is3D ? -1 : 0 }, smax[3] = { 1, 1, mg->D.is3D ? 1 : 0 };
  const int p7[7][3] = { {0,0,0}, {-1,0,0}, {1,0,0}, {0,-1,0}, {0,1,0}, {0,0,-1}, {0,0,1} };
  const int V[3] = { 1, 1, 1 };
  for (int uz = V[2] * 2 + smin[2]; uz <= V[2] * 2 + smax[2]; uz++)
  for (int uy = V[1] * 2 + smin[1]; uy <= V[1] * 2 + smax[1]; uy++)
  for (int ux = V[0] * 2 + smin[0]; ux <= V[0] * 2 + smax[0]; ux++) {
    const int U[3] = { ux, uy, uz };
    for (int i = 0; i < 1 + 2 * dim; i++) {
      const int W[3] = { U[0] + p7[i][0], U[1] + p7[i][1], U[2] + p7[i][2] };
      for (int nz = W[2] / 2; nz <= (W[2] + 1) / 2; nz++)
      for (int ny = W[1] / 2; ny <= (W[1] + 1) / 2; ny++)
      for (int nx = W[0] / 2; nx <= (W[0] + 1) / 2; nx++) {
        const int sidx = nx + 3 * ny + 9 * nz;
        if (sidx >= 13) {
          P p;
          p.N[0] = nx - 1; p.N[1] = ny - 1; p.N[2] = nz - 1;
          for (int d = 0; d < 3; d++) { p.U[d] = U[d] - V[d] * 2; p.W[d] = W[d] - V[d] * 2; }
          p.sc = sidx - 13; p.sf = (i + 1) / 2; p.inU = (i % 2 == 0);
          p.rw = 1.0f / (float)(1 << ((U[0] % 2) + (U[1] % 2) + (U[2] % 2)));
          p.iw = 1.0f / (float)(1 << ((W[0] % 2) + (W[1] % 2) + (W[2] % 2)));
          paths.push_back(p);
        }
      }
    }
  }
  auto pathLess = [](const P& a, const P& b) {
    const int ka = (a.U[0] + 1) + 3 * (a.U[1] + 1) + 9 * (a.U[2] + 1), kb = (b.U[0] + 1) + 3 * (b.U[1] + 1) + 9 * (b.U[2] + 1);
    if (a.sc == b.sc) return ka < kb;
    return a.sc < b.sc;
  };
  std::sort(paths.begin(), paths.end(), pathLess);
  std::vector<MgPath> hp(paths.size());
  for (size_t q = 0; q < paths.size(); q++) {
    const P& p = paths[q]; MgPath& o = hp[q];
    o.Nx = p.N[0]; o.Ny = p.N[1]; o.Nz = p.N[2]; o.Ux = p.U[0]; o.Uy = p.U[1]; o.Uz = p.U[2];
    o.Wx = p.W[0]; o.Wy = p.W[1]; o.Wz = p.W[2]; o.sc = p.sc; o.sf = p.sf; o.inU = p.inU ? 1 : 0; o.rw = p.rw; o.iw = p.iw;
  }
  mg->npaths = (int)hp.size();
  CUDA_CHECK(cudaMalloc(&mg->dPaths, sizeof(MgPath) * hp.size()));
  CUDA_CHECK(cudaMemcpy(mg->dPaths, hp.data(), sizeof(MgPath) * hp.size(), cudaMemcpyHostToDevice));
}

// stride (in floats) between operator planes: every plane is split over the GPUs by z like the vectors, so on a distributed
// solver a plane covers a whole number of allocation granules
static long long planeStride(const Solver* s, long long n, size_t elem = sizeof(float)) {
  if (!(s->comm && !comm_emulated(s->comm))) return n;
  const long long unit = (long long)(comm_granularity(s->comm) / elem);
  return (n + unit - 1) / unit * unit;
}

static void setLayout(LvlD& L, int sx, int sy, int sz, bool is3D, bool blocked) {
  L.sx = sx; L.sy = sy; L.sz = sz; L.n = sx * sy * sz;
  L.perm = blocked ? 1 : 0;
  L.NC = is3D ? 8 : 4; L.zoff = is3D ? 1 : 0;
  L.PX = (sx + 1) / 2 + 2; L.PY = (sy + 1) / 2 + 2; L.PZ = is3D ? (sz + 1) / 2 + 2 : 1;
  const long long na = blocked ? (long long)L.PZ * L.NC * L.PY * L.PX : (long long)L.n;
  mfAssert(na * 14 < (1LL << 40) && na < (1LL << 31), "GridMg: level too large for 32-bit vertex indices");
  L.na = (int)na; L.as = na;
  L.A = L.x = L.b = L.r = nullptr; L.type = nullptr; L.code = nullptr; L.Ah = nullptr; L.ash = 0;
  const unsigned long long BX = (unsigned long long)(L.PX - 2);
  L.bxMagic = BX > 1 ? (unsigned)(((1ULL << 32) + BX - 1) / BX) : 0u;
}

// byte offsets of the stencil terms of a blocked level, per colour (MgOff1); depends on the operator's storage format
static void buildOffsets(MgHierarchy* mg, int l) {
  const LvlD& L = mg->lv[l];
  if ((int)mg->offs.size() <= l) mg->offs.resize(l + 1);
  MgOff8& O = mg->offs[l];
  memset(&O, 0, sizeof(O));
  const bool is3D = mg->D.is3D != 0;
  const int SS = mg->D.SS;
  const long long PXY = (long long)L.PX * L.PY;
  const bool half = L.Ah != nullptr;
  const long long as = half ? L.ash : L.as;
  for (int c = 0; c < L.NC; c++) {
    const int cx = c & 1, cy = (c >> 1) & 1, cz = (c >> 2) & 1;
    int s = 0;
    for (int dz = is3D ? -1 : 0; dz <= (is3D ? 1 : 0); dz++)
      for (int dy = -1; dy <= 1; dy++)
        for (int dx = -1; dx <= 1; dx++, s++) {
          const int nc = ((cx + dx) & 1) | (((cy + dy) & 1) << 1) | (((cz + dz) & 1) << 2);
          // floor((c + d) / 2) for c + d in -1..2
          const long long dp = (long long)((cz + dz + 2) / 2 - 1) * L.NC * PXY + (long long)(nc - c) * PXY
                             + (long long)((cy + dy + 2) / 2 - 1) * L.PX + ((cx + dx + 2) / 2 - 1);
          O.t[c].xo[s] = 4 * dp;
          // stored entry e of the row at shift d: float planes 4 * (e * as + d); binary16 plane pairs 4 * (mgHPair(e) * as + d) + 2 * mgHHalf(e)
          const long long e = (s < SS) ? (SS - 1 - s) : (s - SS + 1), d = (s < SS) ? dp : 0;
          O.t[c].ao[s] = half ? 4 * (mgHPair((int)e) * as + d) + 2 * mgHHalf((int)e) : 4 * (e * as + d);
        }
  }
}

MgHierarchy* mg_create(Solver* s) {
  mfAssert(s->N < (1LL << 31), "GridMg: grid too large for 32-bit vertex indices");
  MgHierarchy* mg = new MgHierarchy();
  mg->s = s; mg->code0 = nullptr; mg->useCode = false; mg->pendingSolo = false; mg->aSet = false; mg->hasTrivial = false; mg->nonZeroStencilSum = false;
  mg->coarsestAcc = 1E-8f; mg->trivialScale = 1E-6f; mg->pre = 1; mg->post = 1; mg->dPaths = nullptr; mg->npaths = 0; mg->usedFallback = 0;
  mg->half1 = nullptr; mg->ash1 = 0; mg->nspecial1 = 0; mg->special1 = nullptr; mg->special1Local = nullptr;
  mg->tailCluster = 0; mg->tailBar = nullptr; mg->tailBase = 0; mg->tailErrHost = nullptr; mg->tailErrDev = nullptr; mg->tailFirst = -1; mg->tailOffs = nullptr;
  mg->bottomOn = false; mg->bottomOff = nullptr; mg->bottomSmem = 0; mg->bottomVec = 0;
  mg->begun = false; mg->begunDst = nullptr; mg->begunRhs = nullptr; mg->begunR0 = nullptr;
  const bool is3D = s->nz > 1;
  mg->D.is3D = is3D; mg->D.dim = is3D ? 3 : 2; mg->D.SS = is3D ? 14 : 5; mg->D.SS0 = is3D ? 4 : 3;
  mg->D.zmin = is3D ? -1 : 0; mg->D.zmax = is3D ? 1 : 0;
  int sx = s->nx, sy = s->ny, sz = s->nz;
  long long n = (long long)sx * sy * sz;
  auto addLevel = [&](int l) {
    LvlD L; setLayout(L, sx, sy, sz, is3D, l > 0);
    // every array is zero-filled by galloc.  Operator planes are separate components, so that each plane is split over the
    // GPUs by z like the vectors; the plane stride is padded to the allocation granularity when the level is distributed.
    // Hosting follows the sweep partition: rank r hosts the granules nearest to its z pairs, rank 0 all of a solo level.
    mg->lv.push_back(L);
    const int P = s->nranks;
    std::vector<int> ps(P + 1, 0);
    std::vector<long long> es(P + 1, 0);
    if (l > 0) {
      const int BZ = is3D ? L.PZ - 2 : 1;
      const long long pair = (long long)L.NC * L.PY * L.PX;
      const bool solo = lvlSolo(mg, l);
      for (int r = 0; r <= P; r++) ps[r] = solo ? (r == 0 ? 0 : BZ) : (int)((long long)BZ * r / P);
      if (!solo && s->comm && !comm_emulated(s->comm)) {
        // move the pair boundaries onto the allocation granules the vectors are split at (4-byte entries)
        const long long gran = (long long)(comm_granularity(s->comm) / sizeof(float));
        for (int r = 1; r < P; r++) {
          const long long e = (ps[r] + L.zoff) * pair;                      // first entry of rank r's share
          const long long eg = (e + gran / 2) / gran * gran;               // start of the granule rank r hosts from
          int q = (int)((eg + pair - 1) / pair) - L.zoff;                   // first pair starting at or behind it
          if (q < ps[r - 1]) q = ps[r - 1];
          if (q > BZ) q = BZ;
          ps[r] = q;
        }
      }
      for (int r = 0; r <= P; r++) es[r] = (r == 0) ? 0 : (r == P ? (solo ? 0 : L.na) : (ps[r] + L.zoff) * pair);
      if (solo) for (int r = 1; r <= P; r++) es[r] = L.na;
    } else {
      for (int r = 0; r <= P; r++) es[r] = (long long)L.sx * L.sy * s->zsplit[r];
    }
    mg->pairSplit.push_back(ps);
    mg->lv.pop_back();
    L.r = (float*)s->galloc(sizeof(float) * (size_t)L.na, 1, 0, es.data(), 4);
    L.type = (unsigned char*)s->galloc((size_t)L.na, 1, 0, es.data(), 1);
    if (l > 0) {
      L.as = planeStride(s, L.na);
      L.A = (float*)s->galloc(sizeof(float) * (size_t)L.na, mg->D.SS, sizeof(float) * (size_t)L.as, es.data(), 4);
      L.x = (float*)s->galloc(sizeof(float) * (size_t)L.na, 1, 0, es.data(), 4);
      L.b = (float*)s->galloc(sizeof(float) * (size_t)L.na, 1, 0, es.data(), 4);
      if (l == 1 && is3D && !getenv("MF_MG_NOHALF")) {
        // seven planes of half2 elements (LvlD::Ah): partitioned like the float planes
        mg->ash1 = L.as;
        mg->half1 = (__half*)s->galloc(sizeof(float) * (size_t)L.na, (mg->D.SS + 1) / 2, sizeof(float) * (size_t)mg->ash1, es.data(), 4);
        mg->special1 = (int*)s->galloc(sizeof(int) * (8 + 8 * MG_SPECIAL_CAP));
        CUDA_CHECK(cudaMalloc(&mg->special1Local, sizeof(int) * (8 + 8 * MG_SPECIAL_CAP)));
      }
    }                                                      // level 0: A (float planes) or code0 are allocated by mg_set_a
    mg->lv.push_back(L);
  };
  addLevel(0);
  for (int l = 1; l <= 100; l++) {          // multigrid.cpp:256-276
    if (sx <= 5 && sy <= 5 && sz <= 5) break;
    if (n <= 1000) break;
    sx = (sx + 2) / 2; sy = (sy + 2) / 2; sz = (sz + 2) / 2;
    n = (long long)sx * sy * sz;
    addLevel(l);
  }
  mfAssert(mg->lv.back().n <= 1024, "GridMg: coarsest level has more than 1024 vertices");
  mg->dFlags = (int*)s->galloc(sizeof(int) * 8);
  buildPaths(mg);
  return mg;
}

void mg_destroy(MgHierarchy* mg) {
  if (!mg) return;
  cudaStreamSynchronize(mg->s->stream);
  for (size_t l = 0; l < mg->lv.size(); l++) {
    LvlD& L = mg->lv[l];
    Solver* s = mg->s;
    if (L.A) s->gfree(L.A);
    s->gfree(L.r); s->gfree(L.type);
    if (l > 0) { s->gfree(L.x); s->gfree(L.b); }
  }
  if (mg->code0) mg->s->gfree(mg->code0);
  if (mg->half1) mg->s->gfree(mg->half1);
  if (mg->special1) mg->s->gfree(mg->special1);
  if (mg->special1Local) cudaFree(mg->special1Local);
  mg->s->gfree(mg->dFlags); cudaFree(mg->dPaths);
  if (mg->tailBar) cudaFree(mg->tailBar);
  if (mg->tailOffs) cudaFree(mg->tailOffs);
  if (mg->bottomOff) cudaFree(mg->bottomOff);
  if (mg->tailErrHost) cudaFreeHost(mg->tailErrHost);
  delete mg;
}

bool mg_is_a_set(const MgHierarchy* mg) { return mg->aSet; }
void mg_set_coarsest_accuracy(MgHierarchy* mg, float acc) { mg->coarsestAcc = acc; }
void mg_set_smoothing(MgHierarchy* mg, int pre, int post) { mg->pre = pre; mg->post = post; }
int mg_num_levels(const MgHierarchy* mg) { return (int)mg->lv.size(); }
int mg_used_host_fallback(const MgHierarchy* mg) { return mg->usedFallback; }
int mg_level_format(const MgHierarchy* mg, int l) { return (l == 0 ? mg->useCode : mg->lv[l].Ah != nullptr) ? 2 : 4; }
int mg_level_size(const MgHierarchy* mg, int l, int* sx, int* sy, int* sz) {
  if (l < 0 || l >= (int)mg->lv.size()) return 0;
  *sx = mg->lv[l].sx; *sy = mg->lv[l].sy; *sz = mg->lv[l].sz; return mg->lv[l].n;
}

// The order-dependent branch of coarse-vertex selection (GridMg::genCoarseGrid, multigrid.cpp:520-578), on the host; only taken
// when the order-independent device sweeps leave fine vertices with >= 2 free interpolation vertices (tests: solve_thin_mg).
// The result must be the reference's, so the visiting order is the specification: always the fine vertex with the fewest free
// interpolation vertices, and among equals the one whose count changed (or was set) most recently.
//
// StampedStacks gives that order without the reference's linked lists: one stack per count; (re)filing a vertex pushes a fresh
// (vertex, stamp) entry and bumps the vertex' stamp, which turns every older entry of it into a tombstone that pop() skips.
namespace {
class StampedStacks {
 public:
  StampedStacks(int vertices, int counts) : stacks_(counts), count_(vertices, 0), stamp_(vertices, 0), live_(0) {}
  int countOf(int v) const { return count_[v]; }            // 0 = not filed
  bool empty() const { return live_ == 0; }
  void file(int v, int count) {                             // count >= 1: (re)file; count == 0: withdraw
    if (count_[v] == 0 && count > 0) live_++;
    if (count_[v] > 0 && count == 0) live_--;
    count_[v] = count; stamp_[v]++;
    if (count > 0) stacks_[count].push_back({ v, stamp_[v] });
  }
  int pop() {                                               // the vertex to visit next; it is withdrawn
    for (size_t c = 1; c < stacks_.size(); c++) {
      std::vector<Item>& st = stacks_[c];
      while (!st.empty()) {
        const Item it = st.back(); st.pop_back();
        if (stamp_[it.v] == it.stamp && count_[it.v] == (int)c) { count_[it.v] = 0; stamp_[it.v]++; live_--; return it.v; }
      }
    }
    return -1;
  }
 private:
  struct Item { int v; unsigned stamp; };
  std::vector<std::vector<Item>> stacks_;
  std::vector<int> count_; std::vector<unsigned> stamp_;
  int live_;
};
struct Box3 { int lo[3], hi[3]; };      // inclusive index box
template <class F> void forBox(const Box3& b, F f) {
  for (int z = b.lo[2]; z <= b.hi[2]; z++) for (int y = b.lo[1]; y <= b.hi[1]; y++) for (int x = b.lo[0]; x <= b.hi[0]; x++) f(x, y, z);
}
}

static void coarsenOnHost(MgHierarchy* mg, int l) {
  Solver* s = mg->s;
  LvlD& F = mg->lv[l - 1]; LvlD& C = mg->lv[l];
  const int fdim[3] = { F.sx, F.sy, F.sz }, cdim[3] = { C.sx, C.sy, C.sz };
  std::vector<unsigned char> fineStored(F.na), coarseStored(C.na, vtInactive);
  CUDA_CHECK(cudaMemcpyAsync(fineStored.data(), F.type, F.na, cudaMemcpyDeviceToHost, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  auto fineId = [&](int x, int y, int z) { return x + fdim[0] * (y + fdim[1] * z); };           // natural order = the reference's vertex ids
  auto coarseId = [&](int x, int y, int z) { return x + cdim[0] * (y + cdim[1] * z); };
  enum Pick : unsigned char { Open, Chosen, Dropped };
  std::vector<unsigned char> pick(C.n, Open);
  // file every active fine vertex under the number of coarse vertices it interpolates from (1, 2, 4 or 8), in id order
  StampedStacks queue(F.n, mg->D.is3D ? 9 : 5);
  forBox(Box3{ { 0, 0, 0 }, { fdim[0] - 1, fdim[1] - 1, fdim[2] - 1 } }, [&](int x, int y, int z) {
    if (fineStored[mgIdx(F, x, y, z)] != vtInactive) queue.file(fineId(x, y, z), 1 << ((x & 1) + (y & 1) + (z & 1)));
  });
  while (!queue.empty()) {
    const int v = queue.pop();
    const int vx = v % fdim[0], vy = (v / fdim[0]) % fdim[1], vz = v / (fdim[0] * fdim[1]);
    bool haveOne = false;
    // the coarse vertices v interpolates from: the first open one is chosen, further open ones are dropped
    forBox(Box3{ { vx / 2, vy / 2, vz / 2 }, { (vx + 1) / 2, (vy + 1) / 2, (vz + 1) / 2 } }, [&](int ix, int iy, int iz) {
      unsigned char& state = pick[coarseId(ix, iy, iz)];
      if (state != Open) return;
      state = haveOne ? Dropped : Chosen;
      haveOne = true;
      // either way it is no longer open for the fine vertices it restricts from: each loses one free interpolation vertex
      Box3 fineNbrs;
      const int ci[3] = { ix, iy, iz };
      for (int d = 0; d < 3; d++) { fineNbrs.lo[d] = std::max(0, 2 * ci[d] - 1); fineNbrs.hi[d] = std::min(fdim[d] - 1, 2 * ci[d] + 1); }
      forBox(fineNbrs, [&](int rx, int ry, int rz) {
        const int r = fineId(rx, ry, rz), left = queue.countOf(r);
        if (left > 0) queue.file(r, left - 1);
      });
    });
  }
  for (int z = 0; z < cdim[2]; z++) for (int y = 0; y < cdim[1]; y++) for (int x = 0; x < cdim[0]; x++)
    coarseStored[mgIdx(C, x, y, z)] = (pick[coarseId(x, y, z)] == Chosen) ? vtActive : vtInactive;     // knActivateCoarseVertices
  const std::vector<unsigned char>& cts = coarseStored;
  // every process computes the same classification; one writes it
  if (s->rankBegin == 0) CUDA_CHECK(cudaMemcpyAsync(C.type, cts.data(), C.na, cudaMemcpyHostToDevice, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  s->rankSync();
  mg->usedFallback++;
}

// per-rank launch of a setup kernel over its storage share of level l; args follow the kernel's leading parameters
#define MG_SETUP(mg, l, kern, threads, ...) do { const bool _solo = lvlSolo(mg, l); beginGroup(mg, _solo); \
  MF_RANKS((mg)->s, _r) { int p0, p1; lvlStore(mg, l, _r, p0, p1); if (p1 > p0) { MF_LAUNCH((mg)->s, kern, cdiv(p1 - p0, threads), threads, __VA_ARGS__, p0, p1); } } \
  endGroup(mg, _solo); } while (0)

static void genCoarseGrid(MgHierarchy* mg, int l) {
  Solver* s = mg->s;
  LvlD& F = mg->lv[l - 1]; LvlD& C = mg->lv[l];
  MG_SETUP(mg, l, k_mg_fill_type, 256, C, (unsigned char)vtFree);
  flushGroups(mg);
  int* flags = mg->dFlags + 4;
  bool pending = false;
  for (int sweep = 0; sweep < 10000; sweep++) {
    s->clearShared(flags, sizeof(int) * 2);
    MG_SETUP(mg, l - 1, k_mg_coarsen_sweep, 256, F, C, flags);
    flushGroups(mg);
    CUDA_CHECK(cudaMemcpyAsync(s->hostInts, flags, sizeof(int) * 2, cudaMemcpyDeviceToHost, s->stream));
    CUDA_CHECK(cudaStreamSynchronize(s->stream));
    pending = s->hostInts[1] != 0;
    if (!s->hostInts[0]) break;
  }
  if (pending) { coarsenOnHost(mg, l); return; }
  MG_SETUP(mg, l, k_mg_coarsen_finalize, 256, C);
  flushGroups(mg);
}

void mg_set_a(MgHierarchy* mg, const float* A0, const float* Ai, const float* Aj, const float* Ak) {
  Solver* s = mg->s;
  LvlD& L0 = mg->lv[0];
  const int maxLevel = (int)mg->lv.size() - 1;
  // try the compact level-0 operator first (any matrix built from flags alone); a single-level hierarchy keeps float
  // planes because the coarsest-level CG reads them
  bool coded = false;
  if (maxLevel > 0) {
    if (!mg->code0) mg->code0 = (unsigned short*)s->galloc(sizeof(unsigned short) * (size_t)L0.n, 1, 0, nullptr, 2);
    s->clearShared(mg->dFlags, sizeof(int) * 4);
    MG_SETUP(mg, 0, k_mg_copy_activate<true>, 256, L0, mg->D, A0, Ai, Aj, Ak, mg->trivialScale, mg->code0, mg->dFlags);
    CUDA_CHECK(cudaMemcpyAsync(s->hostInts, mg->dFlags, sizeof(int) * 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_CHECK(cudaStreamSynchronize(s->stream));
    coded = s->hostInts[2] == 0;
  }
  if (!coded) {
    if (!L0.A) { L0.as = planeStride(s, L0.n); L0.A = (float*)s->galloc(sizeof(float) * (size_t)L0.n, mg->D.SS0, sizeof(float) * (size_t)L0.as); }
    s->clearShared(mg->dFlags, sizeof(int) * 4);
    MG_SETUP(mg, 0, k_mg_copy_activate<false>, 256, L0, mg->D, A0, Ai, Aj, Ak, mg->trivialScale, (unsigned short*)nullptr, mg->dFlags);
    CUDA_CHECK(cudaMemcpyAsync(s->hostInts, mg->dFlags, sizeof(int) * 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_CHECK(cudaStreamSynchronize(s->stream));
  }
  mg->useCode = coded;
  L0.code = coded ? mg->code0 : nullptr;
  mg->nonZeroStencilSum = s->hostInts[0] != 0;
  mg->hasTrivial = s->hostInts[1] != 0;
  for (int l = 1; l <= maxLevel; l++) {
    genCoarseGrid(mg, l);
    LvlD& F = mg->lv[l - 1]; LvlD& C = mg->lv[l];
    if (l == 1) {
      if (coded) MG_SETUP(mg, l, k_mg_galerkin1<L0Code>, 128, F, C, mg->D, mg->dPaths, mg->npaths, mg->trivialScale);
      else       MG_SETUP(mg, l, k_mg_galerkin1<L0Float>, 128, F, C, mg->D, mg->dPaths, mg->npaths, mg->trivialScale);
    } else MG_SETUP(mg, l, k_mg_galerkin, 128, F, C, mg->D);
    flushGroups(mg);
    if (l == 1 && mg->half1) {
      // level 1 carries most of the coarse-level traffic (14 planes, every entry read twice per sweep): keep it in binary16
      // when that is exact, so that the whole level stays resident in L2 between the sweeps of a V-cycle
      // Rows that binary16 cannot hold exactly (the rows around the pinned cell of zeroPressureFixing) stay in float: their
      // vertices are marked with a NaN binary16 diagonal (k_mg_mark_special) and take the float planes in the sweeps.  The level's x
      // vector serves as scratch for the per-row flags and is zeroed again afterwards (it must be 0 at padding / inactive entries).
      int* nspecial = mg->dFlags + 6;
      s->clearShared(nspecial, sizeof(int));
      s->clearShared(mg->special1, sizeof(int) * 8);
      C.Ah = nullptr; C.ash = mg->ash1;
      MG_SETUP(mg, l, k_mg_pack_half, 256, C, mg->D, mg->half1, C.x);
      flushGroups(mg);
      MG_SETUP(mg, l, k_mg_mark_special, 256, C, mg->half1, (const float*)C.x, nspecial, mg->special1, mg->special1 + 8);
      flushGroups(mg);
      MG_SETUP(mg, l, k_mg_zero_range, 256, C.x);
      flushGroups(mg);
      CUDA_CHECK(cudaMemcpyAsync(s->hostInts + 6, nspecial, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
      CUDA_CHECK(cudaMemcpyAsync(s->hostInts + 7, mg->special1, sizeof(int) * 8, cudaMemcpyDeviceToHost, s->stream));
      CUDA_CHECK(cudaMemcpyAsync(mg->special1Local, mg->special1, sizeof(int) * (8 + 8 * MG_SPECIAL_CAP), cudaMemcpyDeviceToDevice, s->stream));
      CUDA_CHECK(cudaStreamSynchronize(s->stream));
      mg->nspecial1 = s->hostInts[6];
      bool fits = true;
      for (int c = 0; c < 8; c++) fits = fits && s->hostInts[7 + c] <= MG_SPECIAL_CAP;
      // binary16 when the whole level converts exactly, or when the rows that do not (around the pinned cell of zeroPressureFixing)
      // fit the per-colour lists the sweeps take from the float planes (MF_MG_HALF_EXACT_ONLY=1: only the former)
      if (mg->nspecial1 == 0 || (fits && !getenv("MF_MG_HALF_EXACT_ONLY"))) C.Ah = mg->half1;
    }
    buildOffsets(mg, l);
  }
  MF_LAUNCH_CHECK();
  mg->aSet = true;
  planTail(mg);
  planBottom(mg);
}

// ---- V-cycle launches: every group = the same kernel on every local rank's z range of the level, then the barrier ----
static inline int rowsBlocks(const LvlD& L, int z0, int z1) { return (int)cdiv((long long)L.sy * (z1 - z0), 8); }   // 8 warps per CTA, one row each

template <class OP>
static void smooth0(MgHierarchy* mg, bool reversed, bool xzero) {
  Solver* s = mg->s;
  LvlD& L = mg->lv[0];
  for (int c = 0; c < 2; c++) {
    const int color = reversed ? 1 - c : c;
    beginGroup(mg, false);
    MF_RANKS(s, r) {
      int z0, z1; lvlPlanes(mg, 0, r, z0, z1);
      if (z1 <= z0) continue;
      if (xzero && c == 0) { MF_LAUNCH_PDL(s, (k_mg_smooth0<OP, true>), rowsBlocks(L, z0, z1), 256, L, mg->D, mg->trivialScale, color, z0, z1); }
      else                 { MF_LAUNCH_PDL(s, (k_mg_smooth0<OP, false>), rowsBlocks(L, z0, z1), 256, L, mg->D, mg->trivialScale, color, z0, z1); }
    }
    endGroup(mg, false);
  }
}

static void smoothGS(MgHierarchy* mg, int l, bool reversed, bool xzero, bool firstDone = false) {
  Solver* s = mg->s; s->profTag = l;
  LvlD& L = mg->lv[l];
  if (l == 0) {
    if (mg->useCode) smooth0<L0Code>(mg, reversed, xzero); else smooth0<L0Float>(mg, reversed, xzero);
    return;
  }
  const bool solo = lvlSolo(mg, l);
  for (int c = firstDone ? 1 : 0; c < L.NC; c++) {     // firstDone: colour 0 was folded into the restriction (mgFoldFirst)
    const int color = reversed ? L.NC - 1 - c : c;
    beginGroup(mg, solo);
    const MgOff1& o = mg->offs[l].t[color];
    MF_RANKS(s, r) {
      int lz0, lz1; lvlPairs(mg, l, r, lz0, lz1);
      if (lz1 <= lz0) continue;
      const dim3 gr(cdiv((long long)(L.PX - 2) * (L.PY - 2), 128), lz1 - lz0);
      const MgSpecial sp = { mg->special1Local, mg->special1Local + 8 };
      if (L.Ah && mg->nspecial1) { MF_LAUNCH_PDL(s, (k_mg_smooth<true, MG_FMT_HALF_LISTS>), dim3(gr.x, gr.y + 1), 128, L, o, color, lz0, sp); }
      else if (L.Ah)       { MF_LAUNCH_PDL(s, (k_mg_smooth<true, MG_FMT_HALF>), gr, 128, L, o, color, lz0, sp); }
      else if (mg->D.is3D) { MF_LAUNCH_PDL(s, (k_mg_smooth<true, MG_FMT_FLOAT>), gr, 128, L, o, color, lz0, sp); }
      else                 { MF_LAUNCH_PDL(s, (k_mg_smooth<false, MG_FMT_FLOAT>), gr, 128, L, o, color, lz0, sp); }
    }
    endGroup(mg, solo);
  }
}

static void residual(MgHierarchy* mg, int l) {
  Solver* s = mg->s; s->profTag = l;
  LvlD& L = mg->lv[l];
  const bool solo = lvlSolo(mg, l);
  beginGroup(mg, solo);
  MF_RANKS(s, r) {
    if (l == 0) {
      int z0, z1; lvlPlanes(mg, 0, r, z0, z1);
      if (z1 <= z0) continue;
      if (mg->useCode) { MF_LAUNCH_PDL(s, k_mg_residual0<L0Code>, rowsBlocks(L, z0, z1), 256, L, mg->D, mg->trivialScale, z0, z1); }
      else             { MF_LAUNCH_PDL(s, k_mg_residual0<L0Float>, rowsBlocks(L, z0, z1), 256, L, mg->D, mg->trivialScale, z0, z1); }
    } else {
      int lz0, lz1; lvlPairs(mg, l, r, lz0, lz1);
      if (lz1 <= lz0) continue;
      const dim3 gr(cdiv((long long)(L.PX - 2) * (L.PY - 2), 128), lz1 - lz0, L.NC);
      const MgOff8& O = mg->offs[l];
      const MgSpecial sp = { mg->special1Local, mg->special1Local + 8 };
      if (L.Ah && mg->nspecial1) { MF_LAUNCH_PDL(s, (k_mg_residual<true, MG_FMT_HALF_LISTS>), dim3(gr.x, gr.y + 1, gr.z), 128, L, O, lz0, sp); }
      else if (L.Ah)       { MF_LAUNCH_PDL(s, (k_mg_residual<true, MG_FMT_HALF>), gr, 128, L, O, lz0, sp); }
      else if (mg->D.is3D) { MF_LAUNCH_PDL(s, (k_mg_residual<true, MG_FMT_FLOAT>), gr, 128, L, O, lz0, sp); }
      else                 { MF_LAUNCH_PDL(s, (k_mg_residual<false, MG_FMT_FLOAT>), gr, 128, L, O, lz0, sp); }
    }
  }
  endGroup(mg, solo);
}
// coarse level l+1 <- fine level l; the coarse level's partition drives the launch
static void restrictTo(MgHierarchy* mg, int l, bool fold) {
  Solver* s = mg->s; s->profTag = l;
  LvlD& F = mg->lv[l]; LvlD& C = mg->lv[l + 1];
  const bool solo = lvlSolo(mg, l + 1);
  // A solo coarse level below a partitioned fine level: every rank restricts the coarse planes whose fine planes it owns and
  // WRITES them into the hosting rank's memory (posted NVLink stores) -- rank 0 alone would gather the fine residual of every
  // peer with NVLink loads (measured at N = 2: 56 us instead of 10 us for the level-1 -> 2 restriction).
  const bool spread = solo && s->multi() && !lvlSolo(mg, l) && mg->D.is3D && !getenv("MF_MG_NOSPREAD");
  beginGroup(mg, solo && lvlSolo(mg, l));     // reads the fine residual: the fine level's writers must have arrived
  MF_RANKS(s, r) {
    int z0, z1;
    if (spread) {
      int f0, f1; lvlPlanes(mg, l, r, f0, f1);                 // fine planes of rank r; coarse plane z belongs to the owner of fine plane 2z
      z0 = (f0 + 1) / 2; z1 = (r == s->nranks - 1) ? C.sz : (f1 + 1) / 2;
      if (r == 0) z0 = 0;
      if (z1 > C.sz) z1 = C.sz;
    } else lvlPlanes(mg, l + 1, r, z0, z1);
    if (z1 <= z0) continue;
    const long long nt = (long long)C.sx * C.sy * (z1 - z0);
    if (mg->D.is3D) { MF_LAUNCH_PDL(s, k_mg_restrict<true>, cdiv(nt, 128), 128, F, C, z0, z1, (int)fold); }
    else            { MF_LAUNCH_PDL(s, k_mg_restrict<false>, cdiv(nt, 128), 128, F, C, z0, z1, (int)fold); }
  }
  endGroup(mg, solo && !spread);              // spread: other ranks wrote the solo level -> a real barrier before rank 0 sweeps it
}
// fine level l += P * coarse level l+1; the fine level's partition drives the launch
static void interpAdd(MgHierarchy* mg, int l) {
  Solver* s = mg->s; s->profTag = l;
  LvlD& F = mg->lv[l]; LvlD& C = mg->lv[l + 1];
  const bool solo = lvlSolo(mg, l);
  beginGroup(mg, solo);
  const bool blocked = F.perm && mg->D.is3D && F.NC == 8 && !getenv("MF_MG_INTERP_GENERIC");
  MF_RANKS(s, r) {
    if (blocked) {
      int lz0, lz1; lvlPairs(mg, l, r, lz0, lz1);
      if (lz1 <= lz0) continue;
      if (getenv("MF_MG_INTERP_PERVERTEX")) { MF_LAUNCH_PDL(s, k_mg_interp_add_b, dim3(cdiv((long long)(F.PX - 2) * (F.PY - 2), 128), lz1 - lz0, 8), 128, F, C, lz0); }
      else { MF_LAUNCH_PDL(s, k_mg_interp_add_c, dim3(cdiv((long long)(F.PX - 2) * (F.PY - 2), 128), lz1 - lz0), 128, F, C, lz0); }
      continue;
    }
    int z0, z1; lvlPlanes(mg, l, r, z0, z1);
    if (z1 <= z0) continue;
    const long long nt = (long long)F.sx * F.sy * (z1 - z0);
    MF_LAUNCH_PDL(s, k_mg_interp_add, cdiv(nt, 256), 256, F, C, z0, z1);
  }
  endGroup(mg, solo);
}

// level 0, first sweep: x = 0, smooth colour 0 then colour 1 in one kernel (x -> L0.x)
template <class OP>
static void fusedPresmooth0(MgHierarchy* mg) {
  Solver* s = mg->s;
  LvlD& L = mg->lv[0];
  s->profTag = 0;
  beginGroup(mg, false);
  MF_RANKS(s, r) {
    int z0, z1; lvlPlanes(mg, 0, r, z0, z1);
    if (z1 > z0) { MF_LAUNCH_PDL(s, k_mg_presmooth0<OP>, rowsBlocks(L, z0, z1), 256, L, mg->trivialScale, 0, z0, z1); }
  }
  endGroup(mg, false);
}
// level 0, up leg: x += P x_coarse, smooth colour 1 then colour 0; reads x (in L0.r) and writes dst
template <class OP>
static void fusedUp0(MgHierarchy* mg, float* dst) {
  Solver* s = mg->s;
  LvlD& L = mg->lv[0]; LvlD& C = mg->lv[1];
  s->profTag = 0;
  beginGroup(mg, false);
  MF_RANKS(s, r) {
    int z0, z1; lvlPlanes(mg, 0, r, z0, z1);
    if (z1 <= z0) continue;
    const dim3 gr(cdiv(L.sx, UP_X), cdiv(L.sy, UP_Y), cdiv(z1 - z0, UP_Z));
    MF_LAUNCH_PDL(s, k_mg_up0<OP>, gr, 256, L, C, mg->trivialScale, (const float*)L.r, dst, 1, z0, z1);
  }
  endGroup(mg, false);
}

// ---- z-marching wavefront legs (mg_wave0.cuh): 16-bit operator code, sx % 4 == 0; partitioned runs need a partitioned level 1
static bool waveOk(const MgHierarchy* mg) {
  const Solver* s = mg->s;
  if (!mg->useCode || mg->lv[0].sx % 4 != 0 || getenv("MF_MG_NOWAVE")) return false;
  return !s->multi() || !lvlSolo(mg, 1);
}
// total CTAs = a whole number of waves of the 2 CTAs an SM holds; chunks of at least ~16 planes (pipeline warm-up is 4-6 planes)
static WvSched waveSched(const Solver* s, int tx, int ty, int planes, int& total) {
  const int tiles = tx * ty, slots = 2 * s->numSMs;
  int maxq = planes / 16; if (maxq < 1) maxq = 1;
  const int waves = (tiles + slots - 1) / slots;
  total = waves * slots;
  if (total > tiles * maxq) total = tiles * maxq;
  if (total < tiles) total = tiles;
  WvSched S; S.tx = tx; S.ty = ty;
  S.q = total / tiles;
  S.nq = tiles - (total - S.q * tiles);           // the remaining total - q*tiles tiles get q+1 chunks
  return S;
}
static void waveDown0(MgHierarchy* mg, bool fold, const int* skip = nullptr) {
  Solver* s = mg->s;
  LvlD& F = mg->lv[0]; LvlD& C = mg->lv[1];
  s->profTag = 0;
  beginGroup(mg, false);
  MF_RANKS(s, r) {
    int z0, z1, cz0, cz1; lvlPlanes(mg, 0, r, z0, z1); lvlPlanes(mg, 1, r, cz0, cz1);
    if (z1 <= z0 && cz1 <= cz0) continue;
    // planes whose x this rank writes, united with the planes 2K of the coarse vertices it writes
    int U0 = z0, U1 = z1;
    if (cz1 > cz0) { if (z1 <= z0) { U0 = 2 * cz0; U1 = 2 * cz1 - 1; } else { U0 = U0 < 2 * cz0 ? U0 : 2 * cz0; U1 = U1 > 2 * cz1 - 1 ? U1 : 2 * cz1 - 1; } }
    U0 &= ~1;
    const int tx = (int)cdiv(F.sx + 1, WV_TX), ty = (int)cdiv(F.sy + 1, WD_TY);
    int total; const WvSched S = waveSched(s, tx, ty, U1 - U0, total);
    MF_LAUNCH_PDL(s, k_mg_down0, total, WV_NT, F, C, mg->trivialScale, 0, U0, U1, S, z0, z1, cz0, cz1, (int)fold, skip);
  }
  endGroup(mg, false);
}
// dot: also reduce dst . rhs and update the CG scalars (MgDotBeta); false when that is not possible here (a rank without
// planes, more CTAs than the reduction holds partials for)
static bool waveUp0(MgHierarchy* mg, float* dst, const MgDotBeta* dot) {
  Solver* s = mg->s;
  LvlD& F = mg->lv[0]; LvlD& C = mg->lv[1];
  s->profTag = 0;
  const int tx = (int)cdiv(F.sx, WV_TX), ty = (int)cdiv(F.sy, WU_TY);
  bool withDot = dot != nullptr && !getenv("MF_MG_NODOT");
  for (int r = 0; r < s->nranks && withDot; r++) {
    int z0, z1; lvlPlanes(mg, 0, r, z0, z1);
    int total = 0;
    if (z1 > z0) waveSched(s, tx, ty, z1 - (z0 & ~1), total);
    if (z1 <= z0 || total > RED_MAX_BLOCKS) withDot = false;
  }
  beginGroup(mg, false);
  WvFinBeta fin; memset(&fin, 0, sizeof(fin));
  if (withDot) fin.d = *dot;
  MF_RANKS(s, r) {
    int z0, z1; lvlPlanes(mg, 0, r, z0, z1);
    if (z1 <= z0) continue;
    const int U0 = z0 & ~1, U1 = z1;
    int total; const WvSched S = waveSched(s, tx, ty, U1 - U0, total);
    if (withDot) { MF_LAUNCH_PDL(s, k_mg_up0w<true>, total, WV_NT, F, C, mg->trivialScale, 1, (const float*)F.r, dst, U0, U1, S, z0, z1, fin, redCtx(s, r)); }
    else         { MF_LAUNCH_PDL(s, k_mg_up0w<false>, total, WV_NT, F, C, mg->trivialScale, 1, (const float*)F.r, dst, U0, U1, S, z0, z1, fin, redCtx(s, r)); }
  }
  // several ranks: the cross-rank barrier that ends the group is the one in front of the combining kernel
  if (withDot && s->multi()) { redFinish<1>(s, RedOps<1>{ { 0 } }, fin); mg->pendingSolo = false; }
  else endGroup(mg, false);
  return withDot;
}

// ---- small levels in two persistent kernels (mg_tail.cuh) ---------------------------------------------------------------------
// The tail starts at the first level whose colour sweep is at most ~one vertex per thread of a one-CTA-per-SM grid; every level
// from there to the coarsest one must be a float, colour-blocked 3-D level and (partitioned solvers) be hosted by rank 0 alone.
static void planTail(MgHierarchy* mg) {
  Solver* s = mg->s;
  mg->tailFirst = -1;
  const int maxLevel = (int)mg->lv.size() - 1;
  // opt-in (MF_MG_TAIL=1): measured at 256^3 the persistent kernels need ~2.7-3.2 us per phase (1.25 us of it the grid barrier,
  // the rest one L2 round trip and the 52-operation dependent chain of a vertex) against ~3.4 us per programmatic dependent
  // launch -- 6.83 vs 6.78 ms per step, no gain; kept for the parity test and as the starting point of a shared-memory variant.
  // MF_MG_TAIL=cluster (r1l): the same kernels as one 16-CTA cluster with the hardware barrier: 2.5 us per phase (barrier 0.24 us,
  // release / acquire 0.5 us, the vertex' ~300 dependent instructions 1.8 us), 5.20 vs 5.04 ms per step -- the phases are bound
  // by the instruction chain of one vertex, not by the barrier or the launch
  if (!mg->D.is3D || maxLevel < 2 || !getenv("MF_MG_TAIL") || getenv("MF_MG_NOTAIL")) return;
  // MF_MG_TAIL=cluster: the kernels run as ONE cluster of 16 (8) CTAs with the hardware cluster barrier between the phases;
  // the tail then starts where a colour has at most one vertex per thread of the cluster
  mg->tailCluster = 0;
  if (!strcmp(getenv("MF_MG_TAIL"), "cluster")) {
    int cl = getenv("MF_MG_TAILCL") ? atoi(getenv("MF_MG_TAILCL")) : 16;
    for (; cl >= 2; cl /= 2) {
      cudaError_t e = cudaSuccess;
      if (cl > 8) {
        e = cudaFuncSetAttribute(k_mg_tail<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_mg_tail<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
      }
      int nc = 0;
      if (e == cudaSuccess) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(cl); cfg.blockDim = dim3(MG_TAIL_THREADS);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        e = cudaOccupancyMaxActiveClusters(&nc, k_mg_tail<true>, &cfg);
      }
      if (e == cudaSuccess && nc >= 1) break;
      (void)cudaGetLastError();
    }
    if (cl < 2) return;
    mg->tailCluster = cl;
  }
  const long long cap = (long long)(mg->tailCluster ? mg->tailCluster : s->numSMs) * MG_TAIL_THREADS;
  int tf = -1;
  for (int l = 1; l < maxLevel; l++) {
    const LvlD& L = mg->lv[l];
    if ((long long)(L.PX - 2) * (L.PY - 2) * (L.PZ - 2) <= cap && !L.Ah) { tf = l; break; }
  }
  if (tf < 0 || maxLevel - tf + 1 > MG_TAIL_MAXL) return;
  for (int l = tf; l <= maxLevel; l++) {
    const LvlD& L = mg->lv[l];
    if (!L.perm || L.Ah || L.NC != 8 || (s->multi() && !lvlSolo(mg, l))) return;
  }
  if (!mg->tailBar) {
    CUDA_CHECK(cudaMalloc(&mg->tailBar, sizeof(unsigned)));
    CUDA_CHECK(cudaMemset(mg->tailBar, 0, sizeof(unsigned)));
    CUDA_CHECK(cudaHostAlloc(&mg->tailErrHost, sizeof(int), cudaHostAllocMapped));
    *mg->tailErrHost = 0;
    CUDA_CHECK(cudaHostGetDevicePointer(&mg->tailErrDev, mg->tailErrHost, 0));
    mg->tailBase = 0;
    CUDA_CHECK(cudaMalloc(&mg->tailOffs, sizeof(MgOff8) * MG_TAIL_MAXL));
  }
  // (stream-ordered behind the kernels of an earlier hierarchy that may still read the tables)
  CUDA_CHECK(cudaMemcpyAsync(mg->tailOffs, &mg->offs[tf], sizeof(MgOff8) * (size_t)(maxLevel - tf), cudaMemcpyHostToDevice, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  mg->tailFirst = tf;
}
// ---- the last swept level and the coarsest level in one CTA (mg_bottom.cuh) ----------------------------------------------------
static void planBottom(MgHierarchy* mg) {
  Solver* s = mg->s;
  mg->bottomOn = false;
  const int maxLevel = (int)mg->lv.size() - 1;
  // opt-in (MF_MG_BOTTOM=1): measured at 256^3 (17^3 + 9^3) the one launch takes 156 us against 124 us for the 18 programmatic
  // dependent launches (~2.2 us each) + the CG kernel (85 us) it replaces, 5.13 vs 5.00 ms per step: 24 warps x ~300 instructions
  // per colour phase are ~1 us of ONE SM's issue slots, the residual eight times that -- the level is too much work for one SM and
  // too little for the launches to matter.  Kept for the parity test.
  if (!mg->D.is3D || maxLevel < 2 || !getenv("MF_MG_BOTTOM") || getenv("MF_MG_NOBOTTOM") || mg->tailFirst >= 0) return;
  const LvlD& L = mg->lv[maxLevel - 1]; const LvlD& B = mg->lv[maxLevel];
  const long long nt = (long long)(L.PX - 2) * (L.PY - 2) * (L.PZ - 2);
  if (!L.perm || !B.perm || L.Ah || L.NC != 8 || nt > MG_BOTTOM_THREADS || B.n > MG_BOTTOM_THREADS) return;
  if (s->multi() && !lvlSolo(mg, maxLevel - 1)) return;
  const int vecDoubles = MG_BOTTOM_THREADS + 2 * (B.sx * B.sy + B.sx + 1);
  const size_t smem = mgBottomSmem(L.na, B.na, vecDoubles);
  if (smem > 200 * 1024) return;
  CUDA_CHECK(cudaFuncSetAttribute(k_mg_bottom, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (!mg->bottomOff) CUDA_CHECK(cudaMalloc(&mg->bottomOff, sizeof(MgOff8)));
  // (stream-ordered behind the kernels of an earlier hierarchy that may still read the table)
  CUDA_CHECK(cudaMemcpyAsync(mg->bottomOff, &mg->offs[maxLevel - 1], sizeof(MgOff8), cudaMemcpyHostToDevice, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  mg->bottomSmem = smem; mg->bottomVec = vecDoubles; mg->bottomOn = true;
}
static void bottomLeg(MgHierarchy* mg, bool fold) {
  Solver* s = mg->s;
  const int maxLevel = (int)mg->lv.size() - 1;
  const bool solo = s->multi();
  s->profTag = maxLevel - 1;
  beginGroup(mg, solo);
  if (s->rankBegin == 0) {
    BottomArgs T;
    T.L = mg->lv[maxLevel - 1]; T.B = mg->lv[maxLevel]; T.D = mg->D; T.off = mg->bottomOff; T.accuracy = mg->coarsestAcc; T.fold = fold ? 1 : 0;
    T.nt = (T.L.PX - 2) * (T.L.PY - 2) * (T.L.PZ - 2); T.vecDoubles = mg->bottomVec;
    MF_LAUNCH_PDL_SMEM(s, mg->bottomSmem, k_mg_bottom, 1, MG_BOTTOM_THREADS, T);
  }
  endGroup(mg, solo);
}
void mg_check_errors(const MgHierarchy* mg) {
  mfAssert(!(mg && mg->tailErrHost && *mg->tailErrHost), "GridMg: a grid barrier of the small-level kernel timed out");
}
static void tailLeg(MgHierarchy* mg, bool down, bool fold) {
  Solver* s = mg->s;
  const int maxLevel = (int)mg->lv.size() - 1, tf = mg->tailFirst;
  const bool solo = s->multi();
  s->profTag = tf;
  beginGroup(mg, solo);
  if (s->rankBegin == 0) {
    TailArgs T;
    memset(&T, 0, sizeof(T));
    T.nl = maxLevel - tf + 1; T.fold = fold ? 1 : 0; T.base = mg->tailBase; T.bar = mg->tailBar; T.err = mg->tailErrDev; T.off = mg->tailOffs; T.prefetch = (down && !getenv("MF_MG_TAILNOPF")) ? 1 : 0;
    static unsigned long long* dbgBuf = nullptr;
    if (getenv("MF_MG_TAILDBG") && !dbgBuf) CUDA_CHECK(cudaMalloc(&dbgBuf, 64 * sizeof(unsigned long long)));
    T.dbg = getenv("MF_MG_TAILDBG") ? dbgBuf : nullptr;
    for (int l = tf; l <= maxLevel; l++) T.lv[l - tf] = mg->lv[l];
    const int G = mg->tailCluster ? mg->tailCluster : s->numSMs;
    T.cluster = mg->tailCluster ? 1 : 0;
    if (!mg->tailCluster) {
      if (down) { MF_LAUNCH_PDL(s, k_mg_tail<true>, G, MG_TAIL_THREADS, T); }
      else      { MF_LAUNCH_PDL(s, k_mg_tail<false>, G, MG_TAIL_THREADS, T); }
    } else {
      mf::ProfScope _ps(s, "k_mg_tail");
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(G); cfg.blockDim = dim3(MG_TAIL_THREADS); cfg.stream = s->stream;
      cudaLaunchAttribute at[2];
      at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = G; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[1].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = at; cfg.numAttrs = s->profiling ? 1 : 2;
      CUDA_CHECK(cudaLaunchKernelEx(&cfg, down ? k_mg_tail<true> : k_mg_tail<false>, T));
      s->launches++;
    }
    mg->tailBase += (unsigned)G * (unsigned)mgTailBarriers(T.nl, down, fold);
    if (T.dbg) {     // MF_MG_TAILDBG: per-phase times of this launch (development)
      unsigned long long h[64];
      CUDA_CHECK(cudaStreamSynchronize(s->stream));
      CUDA_CHECK(cudaMemcpy(h, T.dbg, sizeof(h), cudaMemcpyDeviceToHost));
      const int nb = mgTailBarriers(T.nl, down, fold);
      fprintf(stderr, "k_mg_tail<%s> phases (ns):", down ? "down" : "up");
      for (int i = 1; i <= nb && i < 64; i++) fprintf(stderr, " %llu", h[i] - h[i - 1]);
      fprintf(stderr, "\n");
    }
  }
  endGroup(mg, solo);
}

// setRhs + doVCycle(dst) with src == nullptr (multigrid.cpp:426-433, :448-504)
// the V-cycle starts with the level-0 wavefront down leg
static bool startsWithWave(const MgHierarchy* mg) {
  const int maxLevel = (int)mg->lv.size() - 1;
  return mg->aSet && mg->D.is3D && maxLevel > 0 && mg->pre == 1 && mg->post == 1 && !getenv("MF_MG_UNFUSED") && waveOk(mg) && !getenv("MF_MG_DEVSKIP");
}
bool mg_apply_begin(MgHierarchy* mg, float* dst, const float* rhs, const int* skip) {
  // (not under the per-kernel event profile: a skipped launch would enter the kernel's average)
  if (mg->begun || !startsWithWave(mg) || mg->s->profiling || getenv("MF_MG_NOSPECULATE")) return false;
  mg_check_errors(mg);
  const int maxLevel = (int)mg->lv.size() - 1;
  LvlD& L0 = mg->lv[0];
  L0.b = const_cast<float*>(rhs);
  mg->begunR0 = L0.r; mg->begunDst = dst; mg->begunRhs = rhs;
  L0.x = L0.r; L0.r = dst;
  waveDown0(mg, 1 < maxLevel && !getenv("MF_MG_NOFOLD"), skip);
  mg->s->profTag = -1;
  mg->begun = true;
  return true;
}
void mg_apply_abort(MgHierarchy* mg) {
  if (!mg->begun) return;
  LvlD& L0 = mg->lv[0];
  L0.r = mg->begunR0; L0.x = mg->begunDst;
  flushGroups(mg);
  mg->begun = false;
}
bool mg_apply(MgHierarchy* mg, float* dst, const float* rhs, const MgDotBeta* dot) {
  bool dotDone = false;
  Solver* s = mg->s;
  mfAssert(mg->aSet, "GridMg::setRhs Error: A has not been set.");
  mg_check_errors(mg);
  const int maxLevel = (int)mg->lv.size() - 1;
  LvlD& L0 = mg->lv[0];
  const bool begun = mg->begun;        // mg_apply_begin has queued the level-0 down leg of this very cycle
  mg->begun = false;
  if (begun) { mfAssert(dst == mg->begunDst && rhs == mg->begunRhs, "mg_apply: arguments differ from mg_apply_begin's"); L0.r = mg->begunR0; }
  L0.b = const_cast<float*>(rhs);      // knSetRhs' trivial-row scaling is applied on the fly (OP::rhs)
  L0.x = dst;
  if (maxLevel == 0 || mg->pre == 0) {
    MF_RANKS(s, r) { const Span sp = span(s, r); if (sp.n) CUDA_CHECK(cudaMemsetAsync(dst + sp.off, 0, sizeof(float) * (size_t)sp.n, s->stream)); }
    s->rankSync();
  }
  // level 0 of the standard V(1,1) cycle in 3-D uses the fused kernels of mg_fused0.cuh.  The down leg keeps x in the level's
  // own array (L0.r) and puts the residual into dst, which is free until the tiled up-leg kernel writes it out of place.
  // the restriction onto level l also performs the first pre-smoothing sweep of colour 0 there (not on the coarsest level,
  // whose x is the initial guess of the CG solve)
  const bool noFold = getenv("MF_MG_NOFOLD") != nullptr;
  auto foldTo = [&](int l) { return l > 0 && l < maxLevel && mg->pre >= 1 && !noFold; };
  const bool fused0 = mg->D.is3D && maxLevel > 0 && mg->pre == 1 && mg->post == 1 && !getenv("MF_MG_UNFUSED");
  float* const r0keep = L0.r;
  const bool wave = fused0 && waveOk(mg);
  if (fused0) {
    L0.x = r0keep; L0.r = dst;
    if (wave) { if (!begun) waveDown0(mg, foldTo(1)); }   // both colours, residual and restriction in one pass; the residual stays on chip
    else {
      if (mg->useCode) fusedPresmooth0<L0Code>(mg); else fusedPresmooth0<L0Float>(mg);
      residual(mg, 0);
      restrictTo(mg, 0, foldTo(1));
    }
  }
  // (measured and dropped: a single-CTA kernel walking all levels <= 17^3 plus the coarsest CG in one launch, 229 us against
  // ~256 us for the separate PDL launches at 256^3; a thread-block cluster of 8 x 512 threads walking all phases of a 33^3 or
  // 17^3 level with cluster barriers, 5-9 us per phase against ~4.4 us per dependent launch -- too few threads in flight for
  // the latency of a 41-load stencil item)
  // the last swept level and the coarsest one: one CTA with the level in shared memory (mg_bottom.cuh)
  const bool bottom = mg->bottomOn && mg->pre == 1 && mg->post == 1 && maxLevel - 1 >= (fused0 ? 1 : 0) && !getenv("MF_MG_UNFUSED");
  const int lb = bottom ? maxLevel - 1 : maxLevel;
  // levels tail .. maxLevel-1: both legs as one persistent kernel each (mg_tail.cuh)
  const int tail = (mg->tailFirst >= 1 && mg->pre == 1 && mg->post == 1 && !getenv("MF_MG_NOTAIL") && !getenv("MF_MG_UNFUSED")) ? mg->tailFirst : -1;
  // development (tools/vcycle_time.py --levels): MF_MG_DEVSKIP=l leaves out every level >= l -- WRONG results, per-level timing only
  const int devSkip = getenv("MF_MG_DEVSKIP") ? atoi(getenv("MF_MG_DEVSKIP")) : 1000;
  for (int l = fused0 ? 1 : 0; l < lb; l++) {
    if (l >= devSkip) break;
    if (l == tail) { tailLeg(mg, true, foldTo(l)); break; }
    for (int i = 0; i < mg->pre; i++) smoothGS(mg, l, false, l == 0 && i == 0, i == 0 && foldTo(l));
    residual(mg, l);
    restrictTo(mg, l, foldTo(l + 1));
  }
  if (devSkip <= maxLevel) {}
  else if (bottom) bottomLeg(mg, foldTo(lb));
  else {
    const bool solo = s->multi();        // the bottom levels are always handled by rank 0
    beginGroup(mg, solo);
    s->profTag = lb;
    if (s->rankBegin == 0) {
      const LvlD& B = mg->lv[maxLevel];
      const unsigned nt = (unsigned)((B.n + 31) / 32 * 32);
      if (maxLevel == 0) { if (nt <= 768) { MF_LAUNCH_PDL(s, (k_mg_coarse_cg<true, 768>), 1, nt, B, mg->D, mg->coarsestAcc); }
                           else { MF_LAUNCH_PDL(s, (k_mg_coarse_cg<true, 1024>), 1, nt, B, mg->D, mg->coarsestAcc); } }
      else { if (nt <= 768) { MF_LAUNCH_PDL(s, (k_mg_coarse_cg<false, 768>), 1, nt, B, mg->D, mg->coarsestAcc); }
             else { MF_LAUNCH_PDL(s, (k_mg_coarse_cg<false, 1024>), 1, nt, B, mg->D, mg->coarsestAcc); } }
    }
    endGroup(mg, solo);
  }
  if (tail >= 0) tailLeg(mg, false, false);
  for (int l = (tail >= 0 ? tail : lb) - 1; l >= (fused0 ? 1 : 0); l--) {
    if (l >= devSkip) continue;
    interpAdd(mg, l);
    for (int i = 0; i < mg->post; i++) smoothGS(mg, l, true, false);
  }
  if (fused0) {
    L0.r = r0keep; L0.x = dst;
    if (wave) dotDone = waveUp0(mg, dst, dot);
    else if (mg->useCode) fusedUp0<L0Code>(mg, dst); else fusedUp0<L0Float>(mg, dst);
  }
  flushGroups(mg);
  s->profTag = -1;
  // the trailing calcResidual(0)/calcResidualNorm(0) of doVCycle only feed its return value, which the CG driver
  // ignores (conjugategrad.cpp:162-167) -- skipped.
  return dotDone;
}

void mg_download_level(MgHierarchy* mg, int l, unsigned char* type, float* A) {
  Solver* s = mg->s;
  LvlD& L = mg->lv[l];
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  const int xy = L.sx * L.sy;
  std::vector<unsigned char> ts((size_t)L.na);
  CUDA_CHECK(cudaMemcpy(ts.data(), L.type, (size_t)L.na, cudaMemcpyDeviceToHost));
  if (type) for (int v = 0; v < L.n; v++) type[v] = ts[mgIdx(L, v % L.sx, (v % xy) / L.sx, v / xy)];
  if (A) {
    const int ss = (l == 0) ? mg->D.SS0 : mg->D.SS;
    if (l == 0 && mg->useCode) {
      std::vector<unsigned short> code((size_t)L.n);
      CUDA_CHECK(cudaMemcpy(code.data(), mg->code0, sizeof(unsigned short) * code.size(), cudaMemcpyDeviceToHost));
      for (int v = 0; v < L.n; v++) {
        const int c = code[v];
        float d = (float)codeDiag(c); if (c & MGC_TRIVIAL) d = d * mg->trivialScale;
        A[(size_t)v * ss] = d;
        A[(size_t)v * ss + 1] = (c & MGC_XP) ? -1.f : 0.f;
        A[(size_t)v * ss + 2] = (c & MGC_YP) ? -1.f : 0.f;
        if (ss > 3) A[(size_t)v * ss + 3] = (c & MGC_ZP) ? -1.f : 0.f;
      }
    } else {
      std::vector<float> planes((size_t)L.as * ss);
      CUDA_CHECK(cudaMemcpy(planes.data(), L.A, sizeof(float) * planes.size(), cudaMemcpyDeviceToHost));
      for (int v = 0; v < L.n; v++) {
        const int p = mgIdx(L, v % L.sx, (v % xy) / L.sx, v / xy);
        for (int e = 0; e < ss; e++) A[(size_t)v * ss + e] = planes[(size_t)e * L.as + p];
      }
    }
  }
}

} // namespace mf
