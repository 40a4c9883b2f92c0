// mg_kernels.cuh -- device kernels of the multigrid preconditioner (reference: source/multigrid.cpp).
//
// Layout (DESIGN.md "Multigrid layout"):
//  level 0  : natural x-fastest vertex order (x and b alias the CG vectors).  Operator either as four float planes
//             A[d*n + v] (d = 0 diagonal, 1..3 = +x,+y,+z coupling; the reference's mA[0][v*4+d], multigrid.cpp:351-358)
//             or, when every entry is representable, as one 16-bit code per cell (L0Code below): 2 B instead of 16 B.
//  levels>0 : COLOUR-BLOCKED, z-pair major, zero-padded.  The reference's 27-point Gauss-Seidel sweeps 8 colours = parity
//             of (x,y,z) (multigrid.cpp:713-737).  Vertex (x,y,z) of colour c = (x&1)|(y&1)<<1|(z&1)<<2 is stored at
//                 p = (((z>>1)+1)*8 + c)*PY*PX + ((y>>1)+1)*PX + (x>>1)+1,      PX = (sx+1)/2 + 2, ...
//             (2-D: 4 colours, no z padding).  All vertices of one colour in one pair of z planes are contiguous, and a
//             stencil neighbour (dx,dy,dz) of every vertex of a colour sits in ONE other colour block at a constant index
//             shift, so a colour sweep reads its 27 coefficients and 26 neighbour values fully coalesced (the natural order
//             gives stride-2 access).  z-pair major keeps a z-slab of the level contiguous (multi-GPU hosting).  Padding
//             entries and inactive vertices hold A = x = b = r = 0 and type = inactive, so the stencil loops need neither
//             bounds nor type tests on the neighbours: a zero coefficient leaves the running sum bit-unchanged.
//             The symmetric operator keeps 14 (5 in 2-D) planes A[e*na + p] (reference: mA[l][v*14+e], multigrid.cpp:208-218).
// Arithmetic (summation order, float products) is the reference's; only the storage order differs.
#pragma once
#include "common.cuh"
#include <cuda_fp16.h>
#include "reduce.cuh"

namespace mf {

enum : unsigned char { vtInactive = 0, vtActive = 1, vtActiveTrivial = 2, vtRemoved = 3, vtZero = 4, vtFree = 5 };

struct LvlD {
  int sx, sy, sz;
  int n;               // sx*sy*sz vertices
  int perm;            // 0 natural (level 0), 1 colour-blocked
  int PX, PY, PZ;      // blocked: padded half sizes
  int NC, zoff;        // blocked: colours (8 / 4) and z padding offset (1 / 0)
  int na;              // entries per vector: n (natural) or PZ*NC*PY*PX
  long long as;        // stride between operator planes A[e*as + p] (>= na; a whole number of allocation granules when distributed)
  float* A; float* x; float* b; float* r;
  unsigned char* type;
  const unsigned short* code;   // level 0 only: compact operator (or nullptr)
  unsigned bxMagic;    // blocked: ceil(2^32 / (PX-2)), see mgBlockThread
  // blocked levels: the operator in binary16 (else nullptr).  Planes are interleaved in pairs: entry e of the row stored at position
  // p is the half Ah[2 * (mgHPair(e) * ash + p) + mgHHalf(e)], i.e. seven planes of 4-byte (half2) elements with plane stride ash -- the
  // element size of every other array of the level, so that a z-slab partition of the level cuts the binary16 planes at the same
  // allocation granules as the vectors (2-byte planes put half of a rank's boundary granule on its neighbour: +0.6 ms per step at N = 8)
  const __half* Ah; long long ash;
};

// Pairing of the 14 stored entries of a 3-D row into the seven binary16 plane pairs.  Entry e >= 1 of a row is read by ONE other
// vertex, the neighbour in stencil direction 13 + e, i.e. in the sweep of the colour that differs by the parity of that direction.
// Entries whose directions have the same parity are read in the same sweep from overlapping sectors, so they share an element:
// (2,4) [110], (5,7) and (11,13) [111], (6,12) [011], (8,10) [101]; the diagonal pairs with 1 [100], and 3 [010] with 9 [001]
// (those three elements are half used per sweep): 60 B per vertex and sweep instead of 80 B with the pairing (2q, 2q + 1).
__host__ __device__ constexpr int mgHPair(int e) {
  return (e == 0 || e == 1) ? 0 : (e == 2 || e == 4) ? 1 : (e == 5 || e == 7) ? 2 : (e == 11 || e == 13) ? 3 : (e == 6 || e == 12) ? 4 : (e == 8 || e == 10) ? 5 : 6;
}
__host__ __device__ constexpr int mgHHalf(int e) { return (e == 1 || e == 4 || e == 7 || e == 13 || e == 12 || e == 10 || e == 9) ? 1 : 0; }
__host__ __device__ constexpr int mgHEntry(int q, int h) {
  return h == 0 ? (q == 0 ? 0 : q == 1 ? 2 : q == 2 ? 5 : q == 3 ? 11 : q == 4 ? 6 : q == 5 ? 8 : 3)
                : (q == 0 ? 1 : q == 1 ? 4 : q == 2 ? 7 : q == 3 ? 13 : q == 4 ? 12 : q == 5 ? 10 : 9);
}

struct MgDims { int is3D; int dim; int SS; int SS0; int zmin; int zmax; };   // mStencilSize, mStencilSize0, mStencilMin/Max.z

__device__ __host__ __forceinline__ bool mgInGrid(const LvlD& L, int x, int y, int z) {
  return x >= 0 && y >= 0 && z >= 0 && x < L.sx && y < L.sy && z < L.sz;
}
// storage index of vertex (x,y,z); for blocked levels any x in [-2, 2*(PX-2)+1] etc. is addressable (padding)
__device__ __host__ __forceinline__ int mgIdx(const LvlD& L, int x, int y, int z) {
  if (L.perm) {
    const int c = (x & 1) | ((y & 1) << 1) | ((z & 1) << 2);
    return ((((z >> 1) + L.zoff) * L.NC + c) * L.PY + (y >> 1) + 1) * L.PX + (x >> 1) + 1;
  }
  return x + L.sx * (y + L.sy * z);
}
// storage index p -> vertex coordinates; false for padding entries
__device__ __host__ __forceinline__ bool mgDecode(const LvlD& L, int p, int& x, int& y, int& z) {
  if (p >= L.na) return false;
  if (!L.perm) { x = p % L.sx; const int q = p / L.sx; y = q % L.sy; z = q / L.sy; return true; }
  const int lx = p % L.PX - 1; int q = p / L.PX;
  const int ly = q % L.PY - 1; q /= L.PY;
  const int c = q % L.NC; const int lz = q / L.NC - L.zoff;
  x = 2 * lx + (c & 1); y = 2 * ly + ((c >> 1) & 1); z = 2 * lz + ((c >> 2) & 1);
  return lx >= 0 && ly >= 0 && lz >= 0 && x < L.sx && y < L.sy && z < L.sz;
}
// all-zero row (inactive vertex or padding) of a blocked level: the diagonal is 0
__device__ __forceinline__ bool mgRowZero(const LvlD& L, int p) {
  return L.Ah ? (__half_as_ushort(L.Ah[2 * (long long)p]) & 0x7fffu) == 0 : L.A[p] == 0.f;
}
// (binary16 levels mark a vertex whose row -- or a row it reads -- is not exactly representable with a NaN diagonal: such a
//  vertex takes its coefficients from the float planes, see k_mg_mark_special)
__device__ __forceinline__ float mgDiag(const LvlD& L, int p) {
  if (!L.Ah) return L.A[p];
  const float d = __half2float(L.Ah[2 * (long long)p]);
  return d == d ? d : L.A[p];
}
// setup kernels walk the storage range [p0, p1) (a rank's share of the level)
#define MG_P_FROM_THREAD(L) \
  const int p = p0 + blockIdx.x * blockDim.x + threadIdx.x; \
  int vx, vy, vz; \
  if (p >= p1 || !mgDecode((L), p, vx, vy, vz)) return;

// ---- level-0 operator -------------------------------------------------------------------------------------
// 16-bit code: bit0/1/2 coupling to +x/+y/+z (coefficient -1), bit3/4/5 coupling to -x/-y/-z, bits 6-8 the diagonal
// (0..7), bit 9 set when the whole row is the identity (ActiveTrivial, multigrid.cpp:345-347).
// Representable whenever the matrix comes from flags alone (no fractions, no ghost fluid): MakeLaplaceMatrix
// (conjugategrad.h:154-171) only produces A0 in {0..6} and Ai,Aj,Ak in {0,-1}; fixPressure (pressure.cpp:226-245) keeps that.
enum { MGC_XP = 1, MGC_YP = 2, MGC_ZP = 4, MGC_XM = 8, MGC_YM = 16, MGC_ZM = 32, MGC_DIAG_SHIFT = 6, MGC_TRIVIAL = 512 };
__device__ __host__ __forceinline__ int codeDiag(int c) { return (c >> MGC_DIAG_SHIFT) & 7; }

// knCopyA + knActivateVertices + analyzeStencil (multigrid.cpp:321-384).  CODE: write the 16-bit code (found[2] is set when a
// row is not representable); otherwise the float planes.
template <bool CODE>
__global__ void __launch_bounds__(256) k_mg_copy_activate(LvlD L, MgDims D, const float* __restrict__ A0, const float* __restrict__ Ai,
                                                          const float* __restrict__ Aj, const float* __restrict__ Ak, float trivialScale,
                                                          unsigned short* __restrict__ code,
                                                          int* __restrict__ found /* [0]=nonZeroStencilSum, [1]=trivial, [2]=not codable */,
                                                          int p0, int p1) {
  MG_P_FROM_THREAD(L)
  const int v = p;
  float a[7];
  a[0] = A0[v]; a[1] = Ai[v]; a[2] = Aj[v]; a[3] = D.is3D ? Ak[v] : 0.f;
  a[4] = vx != 0 ? Ai[v - 1] : 0.f;
  a[5] = vy != 0 ? Aj[v - L.sx] : 0.f;
  a[6] = (vz != 0 && D.is3D) ? Ak[v - L.sx * L.sy] : 0.f;
  unsigned char t = vtInactive;
  float diag = a[0];
  bool trivial = false;
  if (a[0] != 0.f) {
    t = vtActive;
    float stencilMax = 0.f, stencilSum = 0.f;
    for (int q = 0; q < 7; q++) { stencilSum += a[q]; stencilMax = fmaxf(stencilMax, fabsf(a[q])); }
    if (fabsf(stencilSum / stencilMax) > 1E-6f) found[0] = 1;
    trivial = a[0] == 1.f && a[1] == 0.f && a[2] == 0.f && a[3] == 0.f && a[4] == 0.f && a[5] == 0.f && a[6] == 0.f;
    if (trivial) { t = vtActiveTrivial; diag = a[0] * trivialScale; found[1] = 1; }
  }
  L.type[v] = t;
  if (CODE) {
    bool ok = a[0] >= 0.f && a[0] <= 7.f && a[0] == (float)(int)a[0];
    int c = ok ? ((int)a[0] << MGC_DIAG_SHIFT) : 0;
    // neighbours at the grid boundary: the matching coefficient of an in-range row must be 0 (never read by the stencils)
#pragma unroll
    for (int q = 1; q < 7; q++) {
      if (a[q] == -1.f) c |= (1 << (q - 1));
      else if (a[q] != 0.f) ok = false;
    }
    if ((vx == L.sx - 1 && a[1] != 0.f) || (vy == L.sy - 1 && a[2] != 0.f) || (D.is3D && vz == L.sz - 1 && a[3] != 0.f)) ok = false;
    if (trivial) c |= MGC_TRIVIAL;
    if (!ok) found[2] = 1;
    code[v] = (unsigned short)c;
  } else {
    L.A[v] = diag; L.A[L.as + v] = a[1]; L.A[2 * L.as + v] = a[2];
    if (D.is3D) L.A[3 * L.as + v] = a[3];
  }
}

__global__ void k_mg_fill_type(LvlD L, unsigned char val, int p0, int p1) {
  MG_P_FROM_THREAD(L)
  L.type[p] = val;
}

// one sweep of the order-independent part of genCoarseGrid (multigrid.cpp:520-578): a fine vertex that is still in the
// heap with exactly one free interpolation vertex selects it ('zero').  flags[0] = changed, flags[1] = some vertex still
// has >= 2 free interpolation vertices (the serial algorithm would have to pop it: order-dependent -> host fallback).
__global__ void __launch_bounds__(256) k_mg_coarsen_sweep(LvlD F, LvlD C, int* __restrict__ flags, int p0, int p1) {
  MG_P_FROM_THREAD(F)
  if (F.type[p] == vtInactive) return;
  int cnt = 0, last = -1;
  for (int z = vz / 2; z <= (vz + 1) / 2; z++)
    for (int y = vy / 2; y <= (vy + 1) / 2; y++)
      for (int x = vx / 2; x <= (vx + 1) / 2; x++) {
        const int i = mgIdx(C, x, y, z);
        if (((volatile unsigned char*)C.type)[i] == vtFree) { cnt++; last = i; }
      }
  if (cnt == 1) { C.type[last] = vtZero; flags[0] = 1; }
  else if (cnt >= 2) flags[1] = 1;
}
// knActivateCoarseVertices (multigrid.cpp:507-516)
__global__ void k_mg_coarsen_finalize(LvlD L, int p0, int p1) {
  MG_P_FROM_THREAD(L)
  unsigned char c = L.type[p];
  if (c == vtFree) c = vtRemoved;
  if (c == vtZero) c = vtActive;
  if (c == vtRemoved) c = vtInactive;
  L.type[p] = c;
}

// ---- level-0 operator policies: float planes or compact code ------------------------------------------------
struct L0Float {
  const float* A; const unsigned char* type; long long n; float trivialScale;
  __device__ __forceinline__ L0Float(const LvlD& L, float ts) : A(L.A), type(L.type), n(L.as), trivialScale(ts) {}
  struct Row { unsigned char t; };
  __device__ __forceinline__ Row row(int v) const { Row r; r.t = type[v]; return r; }
  __device__ __forceinline__ bool inactive(const Row& r) const { return r.t == vtInactive; }
  __device__ __forceinline__ float rhs(const Row& r, float b) const { return r.t == vtActiveTrivial ? b * trivialScale : b; }   // knSetRhs :417-424
  __device__ __forceinline__ float diag(const Row&, int v) const { return A[v]; }
  // coefficient of the edge leaving v in direction DIR (0:-x 1:+x 2:-y 3:+y 4:-z 5:+z); nb = neighbour index
  template <int DIR> __device__ __forceinline__ float coef(const Row&, int v, int nb) const {
    return A[(1 + DIR / 2) * n + ((DIR & 1) ? v : nb)];
  }
  template <int DIR> __device__ __forceinline__ float sub(const Row& r, int v, int nb, float sum, float xn) const { return sum - coef<DIR>(r, v, nb) * xn; }
};
struct L0Code {
  const unsigned short* code; float trivialScale;
  __device__ __forceinline__ L0Code(const LvlD& L, float ts) : code(L.code), trivialScale(ts) {}
  struct Row { int c; };
  __device__ __forceinline__ Row row(int v) const { Row r; r.c = code[v]; return r; }
  __device__ __forceinline__ bool inactive(const Row& r) const { return codeDiag(r.c) == 0; }
  __device__ __forceinline__ float rhs(const Row& r, float b) const { return (r.c & MGC_TRIVIAL) ? b * trivialScale : b; }
  // (float)d for d in 0..7 without an int->float conversion: 2^23 + d is exact, subtracting 2^23 leaves d
  __device__ __forceinline__ float diag(const Row& r, int) const {
    const float d = __int_as_float(0x4B000000 | codeDiag(r.c)) - 8388608.0f;
    return (r.c & MGC_TRIVIAL) ? d * trivialScale : d;
  }
  template <int DIR> __device__ __forceinline__ float coef(const Row& r, int, int) const {
    constexpr int bit = (DIR == 0) ? MGC_XM : (DIR == 1) ? MGC_XP : (DIR == 2) ? MGC_YM : (DIR == 3) ? MGC_YP : (DIR == 4) ? MGC_ZM : MGC_ZP;
    return (r.c & bit) ? -1.0f : 0.0f;
  }
  // sum - coef * xn with coef in {-1, 0}: -(-1 * xn) is exactly +xn, and subtracting 0 * xn = +-0 leaves sum unchanged (up to
  // the sign of a zero sum), so one predicated add replaces select + multiply + subtract
  template <int DIR> __device__ __forceinline__ float sub(const Row& r, int, int, float sum, float xn) const {
    constexpr int bit = (DIR == 0) ? MGC_XM : (DIR == 1) ? MGC_XP : (DIR == 2) ? MGC_YM : (DIR == 3) ? MGC_YP : (DIR == 4) ? MGC_ZM : MGC_ZP;
    return (r.c & bit) ? sum + xn : sum;
  }
};

// b - sum_offdiag(A x) in the reference's order (-x,+x,-y,+y,-z,+z; multigrid.cpp:686-689, :750-753)
template <class OP>
__device__ __forceinline__ float mgOffdiag0(const OP& op, const typename OP::Row& row, const LvlD& L, const MgDims& D, const float* __restrict__ x,
                                            int v, int vx, int vy, int vz, float sum) {
  const int Y = L.sx, Z = L.sx * L.sy;
  if (vx > 0)        sum = op.template sub<0>(row, v, v - 1, sum, x[v - 1]);
  if (vx < L.sx - 1) sum = op.template sub<1>(row, v, v + 1, sum, x[v + 1]);
  if (vy > 0)        sum = op.template sub<2>(row, v, v - Y, sum, x[v - Y]);
  if (vy < L.sy - 1) sum = op.template sub<3>(row, v, v + Y, sum, x[v + Y]);
  if (D.is3D) {
    if (vz > 0)        sum = op.template sub<4>(row, v, v - Z, sum, x[v - Z]);
    if (vz < L.sz - 1) sum = op.template sub<5>(row, v, v + Z, sum, x[v + Z]);
  }
  return sum;
}
// knSmoothColor, l == 0: all vertices with (x+y+z)&1 == color (the reference's two colour groups, :721,:724).
// One warp handles one (y,z) row of the z range [z0,z1).  XZERO: x is known to be all zero (first colour of a V-cycle that
// starts from x = 0, doVCycle with src == nullptr): no neighbour loads, and the other colour / inactive cells are written 0
// so that no separate clear of x is needed.
template <class OP, bool XZERO>
__global__ void __launch_bounds__(256) k_mg_smooth0(LvlD L, MgDims D, float trivialScale, int color, int z0, int z1) {
  MF_PDL_PROLOGUE
  const OP op(L, trivialScale);
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5) + z0 * L.sy;
  if (row >= L.sy * z1) return;
  const int vy = row % L.sy, vz = row / L.sy;
  const int par = (vy + vz + color) & 1;
  const int base = row * L.sx;
  if (XZERO) {
    for (int vx = threadIdx.x & 31; vx < L.sx; vx += 32) {
      const int v = base + vx;
      float out = 0.f;
      if (((vx ^ par) & 1) == 0) {
        const typename OP::Row r = op.row(v);
        if (!op.inactive(r)) out = op.rhs(r, L.b[v]) / op.diag(r, v);
      }
      L.x[v] = out;
    }
  } else {
    for (int vx = 2 * (threadIdx.x & 31) + par; vx < L.sx; vx += 64) {
      const int v = base + vx;
      const typename OP::Row r = op.row(v);
      if (op.inactive(r)) continue;
      const float sum = mgOffdiag0(op, r, L, D, L.x, v, vx, vy, vz, op.rhs(r, L.b[v]));
      L.x[v] = sum / op.diag(r, v);
    }
  }
}
// knCalcResidual, l == 0
template <class OP>
__global__ void __launch_bounds__(256) k_mg_residual0(LvlD L, MgDims D, float trivialScale, int z0, int z1) {
  MF_PDL_PROLOGUE
  const OP op(L, trivialScale);
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5) + z0 * L.sy;
  if (row >= L.sy * z1) return;
  const int vy = row % L.sy, vz = row / L.sy;
  const int base = row * L.sx;
  for (int vx = threadIdx.x & 31; vx < L.sx; vx += 32) {
    const int v = base + vx;
    const typename OP::Row r = op.row(v);
    float sum = 0.f;                     // inactive vertices: r = 0 (the restriction adds them unconditionally)
    if (!op.inactive(r)) {
      sum = mgOffdiag0(op, r, L, D, L.x, v, vx, vy, vz, op.rhs(r, L.b[v]));
      sum -= op.diag(r, v) * L.x[v];
    }
    L.r[v] = sum;
  }
}

// ---- levels > 0 (27-point, 14 stored), colour-blocked -----------------------------------------------------
// The sweeps are instruction-issue and latency bound (one vertex = 27 coefficient loads, 27 loads of x, 54 dependent float
// operations), so the address arithmetic is taken out of them: the host tabulates, per colour and stencil term, the byte offsets
// of the coefficient and of the neighbour's x relative to the vertex' own entries (MgOff1, a kernel parameter: the offsets are
// constant-bank operands of one 64-bit add per load).
struct MgOff1 { long long xo[27]; long long ao[27]; };   // terms in the reference's order (dz, dy, dx ascending)
struct MgOff8 { MgOff1 t[8]; };
struct BlkPos { int p; int lx, ly, lz; };
// work item -> storage position among the vertices of colour c: blockIdx.y = z pair (from lz0), t = position in the pair's
// BX x BY block of the colour (t / BX by multiplication: t * BX < 2^32)
__device__ __forceinline__ bool mgBlockThread(const LvlD& L, int c, int lz0, int t, BlkPos& q, int ypairs, int yb) {
  const int BX = L.PX - 2, BY = L.PY - 2;
  if (t >= BX * BY) return false;
  q.ly = (BX == 1) ? t : (int)__umulhi((unsigned)t, L.bxMagic);
  // z pairs in the order first, last, second, ...: on a z-slab partitioned level the two boundary pairs read their neighbours'
  // values over NVLink; scheduled first, that latency hides behind the interior pairs instead of forming the kernel's tail
  q.lx = t - q.ly * BX; q.lz = lz0 + (yb == 0 ? 0 : (yb == 1 ? ypairs - 1 : yb - 1));
  q.p = (((q.lz + L.zoff) * L.NC + c) * L.PY + q.ly + 1) * L.PX + q.lx + 1;
  return true;
}
// b - sum A x over the stencil in the reference's order (multigrid.cpp:691-706, :755-770); ap / xp = the vertex' own entries of
// operator plane 0 and of x.  The lower half of the stencil is read from the neighbours' rows (symmetric storage).
// b - sum A x in two steps: the row's coefficients (the operator is constant between hierarchy builds, so a kernel launched with
// programmatic dependent launch may load them BEFORE griddepcontrol.wait, while its predecessor drains), then the sum.
// binary16: the vertex' own row (terms SS-1 .. NS-1 = stored entries 0 .. SS-1) lies in plane PAIRS at its own position: one
// 4-byte load per two entries (LvlD::Ah); the lower half of the stencil comes from 13 different neighbours' rows, entry by entry.
template <bool IS3D, bool HALF>
__device__ __forceinline__ void mgLoadRow(const char* __restrict__ ap, const MgOff1& o, float (&a)[IS3D ? 27 : 9]) {
  constexpr int NS = IS3D ? 27 : 9, SS = IS3D ? 14 : 5;
  __half2 own[(SS + 1) / 2];
  if (HALF) {
#pragma unroll
    for (int q = 0; q < (SS + 1) / 2; q++) own[q] = *reinterpret_cast<const __half2*>(ap + o.ao[SS - 1 + mgHEntry(q, 0)]);
  }
#pragma unroll
  for (int s = 0; s < NS; s++) {
    if (HALF && s >= SS - 1) { const int e = s - (SS - 1); a[s] = mgHHalf(e) ? __high2float(own[mgHPair(e)]) : __low2float(own[mgHPair(e)]); }
    else a[s] = HALF ? __half2float(*reinterpret_cast<const __half*>(ap + o.ao[s])) : *reinterpret_cast<const float*>(ap + o.ao[s]);
  }
}
template <bool IS3D, bool WITH_CENTER>
__device__ __forceinline__ float mgRowSub(const float (&a)[IS3D ? 27 : 9], const char* xp, const MgOff1& o, float sum) {
  constexpr int NS = IS3D ? 27 : 9, SS = IS3D ? 14 : 5;
#pragma unroll
  for (int s = 0; s < NS; s++) {
    if (s == SS - 1 && !WITH_CENTER) continue;
    sum -= a[s] * *reinterpret_cast<const float*>(xp + o.xo[s]);
  }
  return sum;
}
// a marked vertex of a binary16 level: the same sum from the float planes.  Stored entry e of a row is stencil term 13 + e; the
// lower half of the stencil (s < 13) is entry 13 - s of the neighbour's row, whose position is p + xo[s] / 4.
template <bool WITH_CENTER>
__device__ __forceinline__ float mgStencilSubFloat(const LvlD& L, int p, const MgOff1& o, float sum) {
  const char* xp = reinterpret_cast<const char*>(L.x + p);
#pragma unroll
  for (int s = 0; s < 27; s++) {
    if (s == 13 && !WITH_CENTER) continue;
    const int e = s >= 13 ? s - 13 : 13 - s;
    const long long pos = s >= 13 ? (long long)p : (long long)p + (o.xo[s] >> 2);
    sum -= L.A[e * L.as + pos] * *reinterpret_cast<const float*>(xp + o.xo[s]);
  }
  return sum;
}
template <bool HALF> __device__ __forceinline__ const char* mgRowPtr(const LvlD& L, int p) {
  return HALF ? reinterpret_cast<const char*>(L.Ah + 2 * (long long)p) : reinterpret_cast<const char*>(L.A + p);
}
// Operator format of a blocked level (template parameter HM of the sweeps): float planes, binary16 planes (every entry converts
// exactly), or binary16 planes + per-colour lists of the vertices that keep the float planes (k_mg_mark_special): those are swept by
// ONE extra CTA row of the same launch (blockIdx.y == 0: scheduled first, its chain of dependent loads -- count, position, row --
// hides behind the streaming CTAs instead of forming the tail), the streaming CTAs only skip them (NaN diagonal).
enum { MG_FMT_FLOAT = 0, MG_FMT_HALF = 1, MG_FMT_HALF_LISTS = 2 };
constexpr int MG_SPECIAL_CAP = 1024;        // vertices per colour list
struct MgSpecial { const int* count; const int* list; };     // count[8], list[8][MG_SPECIAL_CAP]: storage positions
// z pair of storage position p
__device__ __forceinline__ int mgPairOf(const LvlD& L, int p) { return p / (L.PX * L.PY) / L.NC - L.zoff; }

// knSmoothColor, l > 0: one vertex of one of the 8 (4) colours (multigrid.cpp:668-711)
template <bool IS3D, int HM>
__global__ void __launch_bounds__(128, 6) k_mg_smooth(LvlD L, const __grid_constant__ MgOff1 o, int c, int lz0, MgSpecial sp) {
  pdlTrigger();
  constexpr bool HALF = HM != MG_FMT_FLOAT;
  const int ypairs = gridDim.y - (HM == MG_FMT_HALF_LISTS ? 1 : 0);
  BlkPos q;
  const bool listRow = HM == MG_FMT_HALF_LISTS && blockIdx.y == 0;
  const bool mine = !listRow && mgBlockThread(L, c, lz0, blockIdx.x * blockDim.x + threadIdx.x, q, ypairs, (int)blockIdx.y - (HM == MG_FMT_HALF_LISTS ? 1 : 0));
  // the operator row before the wait (constant data), x and b behind it
  float a[IS3D ? 27 : 9];
  const char* ap = nullptr;
  if (mine) { ap = mgRowPtr<HALF>(L, q.p); mgLoadRow<IS3D, HALF>(ap, o, a); }
  pdlWait();
  if (listRow) {
    // the listed vertices of this colour in the z pairs [lz0, lz0 + ypairs) of this launch, from the float planes
    if (blockIdx.x != 0) return;
    const int n = min(sp.count[c], MG_SPECIAL_CAP);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int p = sp.list[c * MG_SPECIAL_CAP + i];
      const int lz = mgPairOf(L, p);
      if (lz < lz0 || lz >= lz0 + ypairs) continue;
      const float diag = L.A[p];
      const float sum = mgStencilSubFloat<false>(L, p, o, L.b[p]);
      if (diag != 0.f) L.x[p] = sum / diag;
    }
    return;
  }
  if (!mine) return;
  // inactive vertices and padding have an all-zero row (k_mg_galerkin*); an active row has a non-zero diagonal, or the
  // reference's own division below would not be finite
  // The diagonal is loaded with the rest of the row and only the store is predicated: a vertex is one batch of loads (one trip
  // to L2 / DRAM instead of two) -- the sweeps of the small levels are pure latency.  (128, 6): up to 80 registers, enough for
  // the compiler to issue all 54 loads before the first use (with 32 registers it interleaved them in three dependent rounds).
  const float diag = a[IS3D ? 13 : 4];
  const float sum = mgRowSub<IS3D, false>(a, reinterpret_cast<const char*>(L.x + q.p), o, L.b[q.p]);
  if (HM == MG_FMT_HALF_LISTS) { if (diag != 0.f && diag == diag) L.x[q.p] = sum / diag; }      // NaN: a listed vertex
  else if (diag != 0.f) L.x[q.p] = sum / diag;
}
// knCalcResidual, l > 0; blockIdx.z = colour
template <bool IS3D, int HM, int C>
__device__ __forceinline__ void mgResidualItem(const LvlD& L, const MgOff8& O, int lz0, const MgSpecial& sp) {
  constexpr bool HALF = HM != MG_FMT_FLOAT;
  const int ypairs = gridDim.y - (HM == MG_FMT_HALF_LISTS ? 1 : 0);
  const bool listRow = HM == MG_FMT_HALF_LISTS && blockIdx.y == 0;
  BlkPos q;
  const bool mine = !listRow && mgBlockThread(L, C, lz0, blockIdx.x * blockDim.x + threadIdx.x, q, ypairs, (int)blockIdx.y - (HM == MG_FMT_HALF_LISTS ? 1 : 0));
  // the operator row before the wait (constant data, see k_mg_smooth)
  float a[IS3D ? 27 : 9];
  if (mine) mgLoadRow<IS3D, HALF>(mgRowPtr<HALF>(L, q.p), O.t[C], a);
  pdlWait();
  if (listRow) {
    if (blockIdx.x != 0) return;
    const int n = min(sp.count[C], MG_SPECIAL_CAP);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int p = sp.list[C * MG_SPECIAL_CAP + i];
      const int lz = mgPairOf(L, p);
      if (lz < lz0 || lz >= lz0 + ypairs) continue;
      L.r[p] = mgStencilSubFloat<true>(L, p, O.t[C], L.b[p]);
    }
    return;
  }
  if (!mine) return;
  // inactive / padding entries: b = 0 and an all-zero row give r = 0, what they hold anyway -- no type test needed
  const float r = mgRowSub<IS3D, true>(a, reinterpret_cast<const char*>(L.x + q.p), O.t[C], L.b[q.p]);
  if (HM == MG_FMT_HALF_LISTS) { const float dg = a[IS3D ? 13 : 4]; if (dg == dg) L.r[q.p] = r; }
  else L.r[q.p] = r;
}
template <bool IS3D, int HM>
__global__ void __launch_bounds__(128, 6) k_mg_residual(LvlD L, const __grid_constant__ MgOff8 O, int lz0, MgSpecial sp) {
  pdlTrigger();                 // (griddepcontrol.wait inside the item, behind the loads of the operator row)
  switch (blockIdx.z) {       // the colour selects compile-time positions in the offset table
    case 0: mgResidualItem<IS3D, HM, 0>(L, O, lz0, sp); break;
    case 1: mgResidualItem<IS3D, HM, 1>(L, O, lz0, sp); break;
    case 2: mgResidualItem<IS3D, HM, 2>(L, O, lz0, sp); break;
    case 3: mgResidualItem<IS3D, HM, 3>(L, O, lz0, sp); break;
    case 4: if (IS3D) mgResidualItem<IS3D, HM, 4>(L, O, lz0, sp); break;
    case 5: if (IS3D) mgResidualItem<IS3D, HM, 5>(L, O, lz0, sp); break;
    case 6: if (IS3D) mgResidualItem<IS3D, HM, 6>(L, O, lz0, sp); break;
    default: if (IS3D) mgResidualItem<IS3D, HM, 7>(L, O, lz0, sp); break;
  }
}

// knRestrict (multigrid.cpp:904-927): b_coarse = R r_fine ; also x_coarse = 0 (knSet, :472).  Work items run over the natural
// coordinates of the coarse level, z range [z0,z1) of the coarse level.  r_fine is 0 at inactive / padding / (level 0:
// guarded) out-of-grid vertices.
// fold: the pre-smoothing sweep of colour 0 that follows on x == 0 is x = b / diag (every stencil term is a * 0) -- done here,
// and the sweep starts at colour 1.
__device__ __forceinline__ float mgFoldFirst(const LvlD& C, int v, int vx, int vy, int vz, float b, bool fold) {
  return (fold && ((vx | vy | vz) & 1) == 0) ? b / mgDiag(C, v) : 0.f;
}
template <bool IS3D>
__device__ __forceinline__ void mgRestrictItem(const LvlD& F, const LvlD& C, int z0, int z1, int t, bool fold) {
  const int vx = t % C.sx; const int u = t / C.sx; const int vy = u % C.sy, vz = z0 + u / C.sy;
  if (vz >= z1) return;
  const int v = mgIdx(C, vx, vy, vz);
  const float diag = mgDiag(C, v);        // loaded with the residuals (one trip to memory), tested after the sum
  float sum = 0.f;
  const float* fr = F.r;
  if (F.perm) {
    // fine vertex (2 vx + dx, 2 vy + dy, 2 vz + dz): colour = parity of (dx, dy, dz), position (vx, vy, vz) of its colour block, one
    // entry / row / z pair back for an offset of -1 -- a constant shift from the position of (2 vx, 2 vy, 2 vz) in colour block 0
    const int PXY = F.PX * F.PY;
    const float* f0 = fr + ((((vz + F.zoff) * F.NC) * F.PY + vy + 1) * F.PX + vx + 1);
#pragma unroll
    for (int dz = IS3D ? -1 : 0; dz <= (IS3D ? 1 : 0); dz++)
#pragma unroll
      for (int dy = -1; dy <= 1; dy++)
#pragma unroll
        for (int dx = -1; dx <= 1; dx++) {
          // the reference clamps the loops to the grid: skipped terms, not zero terms -- identical sums because the
          // first in-range term is added to +0 either way
          const float rw = 1.0f / (float)(1 << ((dx & 1) + (dy & 1) + (dz & 1)));
          const int c = (dx & 1) | ((dy & 1) << 1) | ((dz & 1) << 2);
          const int off = (dz < 0 ? -F.NC * PXY : 0) + c * PXY + (dy < 0 ? -F.PX : 0) + (dx < 0 ? -1 : 0);
          sum += rw * f0[off];
        }
  } else {
    for (int z = max(0, vz * 2 - 1); z <= min(F.sz - 1, vz * 2 + 1); z++)
      for (int y = max(0, vy * 2 - 1); y <= min(F.sy - 1, vy * 2 + 1); y++)
        for (int x = max(0, vx * 2 - 1); x <= min(F.sx - 1, vx * 2 + 1); x++) {
          const float rw = 1.0f / (float)(1 << ((x % 2) + (y % 2) + (z % 2)));
          sum += rw * fr[mgIdx(F, x, y, z)];
        }
  }
  if (diag == 0.f) { C.x[v] = 0.f; return; }           // inactive coarse vertex (all-zero row): b stays 0
  C.b[v] = sum;
  C.x[v] = (fold && ((vx | vy | vz) & 1) == 0) ? sum / diag : 0.f;      // mgFoldFirst
}
template <bool IS3D>
__global__ void __launch_bounds__(128) k_mg_restrict(LvlD F, LvlD C, int z0, int z1, int fold) {
  MF_PDL_PROLOGUE
  mgRestrictItem<IS3D>(F, C, z0, z1, blockIdx.x * blockDim.x + threadIdx.x, fold != 0);
}
// knInterpolate + knAddAssign (multigrid.cpp:934-954, :445-446): x_fine += P x_coarse.  Work items run over the natural
// coordinates of the fine level, z range [z0,z1) of the fine level; x_coarse is 0 at inactive / padding vertices.
__device__ __forceinline__ void mgInterpAddItem(const LvlD& F, const LvlD& C, int z0, int z1, int t) {
  const int vx = t % F.sx; const int u = t / F.sx; const int vy = u % F.sy, vz = z0 + u / F.sy;
  if (vz >= z1) return;
  const int v = mgIdx(F, vx, vy, vz);
  const bool off = F.perm ? mgRowZero(F, v) : (F.type[v] == vtInactive);     // tested after the loads (one trip to memory)
  const float xo = F.x[v];
  float sum = 0.f;
  const float* cx = C.x;
  for (int z = vz / 2; z <= (vz + 1) / 2; z++)
    for (int y = vy / 2; y <= (vy + 1) / 2; y++)
      for (int x = vx / 2; x <= (vx + 1) / 2; x++) sum += cx[mgIdx(C, x, y, z)];
  const float iw = 1.0f / (float)(1 << ((vx % 2) + (vy % 2) + (vz % 2)));
  if (!off) F.x[v] = xo + iw * sum;
}
__global__ void __launch_bounds__(256) k_mg_interp_add(LvlD F, LvlD C, int z0, int z1) {
  MF_PDL_PROLOGUE
  mgInterpAddItem(F, C, z0, z1, blockIdx.x * blockDim.x + threadIdx.x);
}

// The same for a colour-blocked fine level, without divisions: a work item is a storage position of colour CF (blockIdx.z) in z
// pair blockIdx.y, so the vertex is (2 lx + cx, 2 ly + cy, 2 lz + cz) and the coarse vertices it interpolates from are
// (lx .. lx + cx, ly .. ly + cy, lz .. lz + cz): compile-time loop bounds, same summation order (z, y, x ascending).
template <int CF>
__device__ __forceinline__ void mgInterpAddBlockedItem(const LvlD& F, const LvlD& C, int lz0) {
  BlkPos q;
  if (!mgBlockThread(F, CF, lz0, blockIdx.x * blockDim.x + threadIdx.x, q, gridDim.y, blockIdx.y)) return;
  constexpr int cx = CF & 1, cy = (CF >> 1) & 1, cz = (CF >> 2) & 1;
  const bool off = mgRowZero(F, q.p);
  const float xo = F.x[q.p];
  float sum = 0.f;
#pragma unroll
  for (int dz = 0; dz <= cz; dz++)
#pragma unroll
    for (int dy = 0; dy <= cy; dy++)
#pragma unroll
      for (int dx = 0; dx <= cx; dx++) sum += C.x[mgIdx(C, q.lx + dx, q.ly + dy, q.lz + dz)];
  const float iw = 1.0f / (float)(1 << (cx + cy + cz));
  if (!off) F.x[q.p] = xo + iw * sum;
}
__global__ void __launch_bounds__(128) k_mg_interp_add_b(LvlD F, LvlD C, int lz0) {
  MF_PDL_PROLOGUE
  switch (blockIdx.z) {
    case 0: mgInterpAddBlockedItem<0>(F, C, lz0); break;
    case 1: mgInterpAddBlockedItem<1>(F, C, lz0); break;
    case 2: mgInterpAddBlockedItem<2>(F, C, lz0); break;
    case 3: mgInterpAddBlockedItem<3>(F, C, lz0); break;
    case 4: mgInterpAddBlockedItem<4>(F, C, lz0); break;
    case 5: mgInterpAddBlockedItem<5>(F, C, lz0); break;
    case 6: mgInterpAddBlockedItem<6>(F, C, lz0); break;
    default: mgInterpAddBlockedItem<7>(F, C, lz0); break;
  }
}

// The same with one work item per 2 x 2 x 2 CELL of fine vertices: the eight vertices (2 lx + cx, 2 ly + cy, 2 lz + cz) sit at the
// SAME position of the eight colour blocks and interpolate from the eight coarse vertices (lx .. lx + 1, ly .. ly + 1, lz .. lz + 1),
// which are loaded once (8 loads instead of 27 for the eight vertices; the per-vertex sums keep the order z, y, x ascending).
__global__ void __launch_bounds__(128) k_mg_interp_add_c(LvlD F, LvlD C, int lz0) {
  pdlTrigger();
  BlkPos q;
  const bool mine = mgBlockThread(F, 0, lz0, blockIdx.x * blockDim.x + threadIdx.x, q, gridDim.y, blockIdx.y);
  const int PXY = F.PX * F.PY;
  bool off[8];                          // all-zero rows: read from the (constant) operator before the wait
  if (mine) {
#pragma unroll
    for (int c = 0; c < 8; c++) off[c] = mgRowZero(F, q.p + c * PXY);
  }
  pdlWait();
  if (!mine) return;
  float cv[2][2][2];
#pragma unroll
  for (int dz = 0; dz < 2; dz++)
#pragma unroll
    for (int dy = 0; dy < 2; dy++)
#pragma unroll
      for (int dx = 0; dx < 2; dx++) cv[dz][dy][dx] = C.x[mgIdx(C, q.lx + dx, q.ly + dy, q.lz + dz)];
  float xo[8];
#pragma unroll
  for (int c = 0; c < 8; c++) xo[c] = F.x[q.p + c * PXY];
#pragma unroll
  for (int c = 0; c < 8; c++) {
    const int cx = c & 1, cy = (c >> 1) & 1, cz = (c >> 2) & 1;
    float sum = 0.f;
#pragma unroll
    for (int dz = 0; dz <= cz; dz++)
#pragma unroll
      for (int dy = 0; dy <= cy; dy++)
#pragma unroll
        for (int dx = 0; dx <= cx; dx++) sum += cv[dz][dy][dx];
    const float iw = 1.0f / (float)(1 << (cx + cy + cz));
    if (!off[c]) F.x[q.p + c * PXY] = xo[c] + iw * sum;
  }
}

// ---- Galerkin coarse operators (setup) ---------------------------------------------------------------------
// precomputed coarsening paths for level 1 (multigrid.cpp:286-318)
struct MgPath { signed char Nx, Ny, Nz, Ux, Uy, Uz, Wx, Wy, Wz; signed char sc, sf, inU; float rw, iw; };

// level-0 coefficient by stencil slot sf (0 diagonal, 1..3 = +x,+y,+z) at storage vertex v
template <class OP> __device__ __forceinline__ float mgCoef0(const OP& op, int sf, int v) {
  const typename OP::Row r = op.row(v);
  if (sf == 0) return op.diag(r, v);
  if (sf == 1) return op.template coef<1>(r, v, v);
  if (sf == 2) return op.template coef<3>(r, v, v);
  return op.template coef<5>(r, v, v);
}
// knGenCoarseGridOperator, l == 1 (multigrid.cpp:594-614)
template <class OP>
__global__ void __launch_bounds__(128) k_mg_galerkin1(LvlD F, LvlD C, MgDims D, const MgPath* __restrict__ paths, int npaths, float trivialScale,
                                                      int p0, int p1) {
  MG_P_FROM_THREAD(C)
  const int v = p;
  if (C.type[v] == vtInactive) { for (int e = 0; e < D.SS; e++) C.A[e * C.as + v] = 0.f; return; }
  const OP op(F, trivialScale);
  float acc[14];
#pragma unroll
  for (int e = 0; e < 14; e++) acc[e] = 0.f;
  for (int q = 0; q < npaths; q++) {
    const MgPath pt = paths[q];
    const int Nx = vx + pt.Nx, Ny = vy + pt.Ny, Nz = vz + pt.Nz;
    if (!mgInGrid(C, Nx, Ny, Nz) || C.type[mgIdx(C, Nx, Ny, Nz)] == vtInactive) continue;
    const int Ux = vx * 2 + pt.Ux, Uy = vy * 2 + pt.Uy, Uz = vz * 2 + pt.Uz;
    if (!mgInGrid(F, Ux, Uy, Uz)) continue;
    const int u = mgIdx(F, Ux, Uy, Uz);
    if (F.type[u] == vtInactive) continue;
    const int Wx = vx * 2 + pt.Wx, Wy = vy * 2 + pt.Wy, Wz = vz * 2 + pt.Wz;
    if (!mgInGrid(F, Wx, Wy, Wz)) continue;
    const int w = mgIdx(F, Wx, Wy, Wz);
    if (F.type[w] == vtInactive) continue;
    const float a = mgCoef0(op, pt.sf, pt.inU ? u : w);
    const float add = pt.rw * a * pt.iw;
    // static indexing keeps acc[] in registers
#pragma unroll
    for (int e = 0; e < 14; e++) if (e == pt.sc) acc[e] += add;
  }
  for (int e = 0; e < D.SS; e++) C.A[e * C.as + v] = acc[e];
}

// knGenCoarseGridOperator, l > 1 (multigrid.cpp:615-656); F has the 27-point layout
__global__ void __launch_bounds__(128) k_mg_galerkin(LvlD F, LvlD C, MgDims D, int p0, int p1) {
  MG_P_FROM_THREAD(C)
  const int v = p;
  if (C.type[v] == vtInactive) { for (int e = 0; e < D.SS; e++) C.A[e * C.as + v] = 0.f; return; }
  float acc[14];
#pragma unroll
  for (int e = 0; e < 14; e++) acc[e] = 0.f;
  const int SS = D.SS;
  const int smz = D.is3D ? 1 : 0;
  for (int Uz = max(0, vz * 2 - 1); Uz <= min(F.sz - 1, vz * 2 + 1); Uz++)
  for (int Uy = max(0, vy * 2 - 1); Uy <= min(F.sy - 1, vy * 2 + 1); Uy++)
  for (int Ux = max(0, vx * 2 - 1); Ux <= min(F.sx - 1, vx * 2 + 1); Ux++) {
    const int u = mgIdx(F, Ux, Uy, Uz);
    if (F.type[u] == vtInactive) continue;
    const float rw = 1.0f / (float)(1 << ((Ux % 2) + (Uy % 2) + (Uz % 2)));
    // (U-1)/2 uses C++ truncating division, as in the reference
    for (int Nz = (Uz - 1) / 2; Nz <= min(C.sz - 1, (Uz + 2) / 2); Nz++)
    for (int Ny = (Uy - 1) / 2; Ny <= min(C.sy - 1, (Uy + 2) / 2); Ny++)
    for (int Nx = (Ux - 1) / 2; Nx <= min(C.sx - 1, (Ux + 2) / 2); Nx++) {
      const int nn = mgIdx(C, Nx, Ny, Nz);
      if (C.type[nn] == vtInactive) continue;
      const int sc = (Nx - vx + 1) + 3 * (Ny - vy + 1) + 9 * (Nz - vz + smz);
      if (sc < SS - 1) continue;
      const int e = sc - SS + 1;
      for (int Wz = max(0, max(Uz - 1, Nz * 2 - 1)); Wz <= min(F.sz - 1, min(Uz + 1, Nz * 2 + 1)); Wz++)
      for (int Wy = max(0, max(Uy - 1, Ny * 2 - 1)); Wy <= min(F.sy - 1, min(Uy + 1, Ny * 2 + 1)); Wy++)
      for (int Wx = max(0, max(Ux - 1, Nx * 2 - 1)); Wx <= min(F.sx - 1, min(Ux + 1, Nx * 2 + 1)); Wx++) {
        const int w = mgIdx(F, Wx, Wy, Wz);
        if (F.type[w] == vtInactive) continue;
        const int sf = (Wx - Ux + 1) + 3 * (Wy - Uy + 1) + 9 * (Wz - Uz + smz);
        const float iw = 1.0f / (float)(1 << ((Wx % 2) + (Wy % 2) + (Wz % 2)));
        float a;
        if (sf < SS) a = F.A[(SS - 1 - sf) * F.as + w];
        else         a = F.A[(sf - SS + 1) * F.as + u];
        const float add = rw * a * iw;
#pragma unroll
        for (int q = 0; q < 14; q++) if (q == e) acc[q] += add;
      }
    }
  }
  for (int e = 0; e < SS; e++) C.A[e * C.as + v] = acc[e];
}

// binary16 copy of a blocked level's operator; *inexact is set when an entry does not survive the round trip.  The level-1
// Galerkin operator of a 0 / -1 / small-integer Laplacian consists of multiples of 1/64 below 32 and converts exactly.
__global__ void __launch_bounds__(256) k_mg_pack_half(LvlD L, MgDims D, __half* __restrict__ Ah, float* __restrict__ rowBad, int p0, int p1) {
  const int p = p0 + blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= p1) return;
  bool bad = false;
  for (int e = 0; e < D.SS; e++) {
    const float a = L.A[e * L.as + p];
    const __half h = __float2half_rn(a);
    bad |= !(__half2float(h) == a);
    Ah[2 * (mgHPair(e) * L.ash + p) + mgHHalf(e)] = h;
  }
  rowBad[p] = bad ? 1.f : 0.f;
}
// A vertex is "special" when its own row or the row of any of its 26 neighbours (a superset of the 13 rows its sweep reads) has an
// entry that binary16 cannot hold -- in practice the rows around the pinned cell of zeroPressureFixing, whose level-0 equation is
// scaled by 1e-6.  Its binary16 diagonal becomes NaN: the sweeps then take that vertex' coefficients from the float planes
// (mgStencilSubFloat), everybody else streams the binary16 copy.  *count = number of special vertices.
__global__ void __launch_bounds__(256) k_mg_mark_special(LvlD L, __half* __restrict__ Ah, const float* __restrict__ rowBad, int* __restrict__ count,
                                                         int* __restrict__ colourCount, int* __restrict__ colourList, int p0, int p1) {
  const int p = p0 + blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= p1) return;
  int x, y, z;
  if (!mgDecode(L, p, x, y, z)) return;
  bool special = false;
  for (int dz = -1; dz <= 1; dz++) for (int dy = -1; dy <= 1; dy++) for (int dx = -1; dx <= 1; dx++) special |= rowBad[mgIdx(L, x + dx, y + dy, z + dz)] != 0.f;
  // (inactive vertices keep their exact all-zero row: the entries that couple a neighbour to them are zero as well, and a NaN
  //  diagonal would make mgRowZero() take them for active)
  if (special && L.A[p] != 0.f) {
    Ah[2 * (long long)p] = __ushort_as_half((unsigned short)0x7e00); atomicAdd(count, 1);
    // per-colour list for the sweeps (vertices of one colour are independent of each other: the order in the list is free)
    const int c = (x & 1) | ((y & 1) << 1) | ((z & 1) << 2);
    const int i = atomicAdd(colourCount + c, 1);
    if (i < MG_SPECIAL_CAP) colourList[c * MG_SPECIAL_CAP + i] = p;
  }
}
__global__ void __launch_bounds__(256) k_mg_zero_range(float* __restrict__ a, int p0, int p1) {
  const int p = p0 + blockIdx.x * blockDim.x + threadIdx.x;
  if (p < p1) a[p] = 0.f;
}

} // namespace mf
