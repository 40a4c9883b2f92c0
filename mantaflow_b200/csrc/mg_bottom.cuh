// mg_bottom.cuh -- the bottom of the V-cycle in ONE CTA (reference: GridMg::doVCycle multigrid.cpp:448-504 for the last swept
// level and the coarsest one: knSmoothColor :668, knCalcResidual :739, knRestrict :904, solveCG :796, knInterpolate :934).
//
// At 256^3 the last swept level has 17^3 vertices and the coarsest one 9^3: a visit is 7 colour sweeps + residual + restriction,
// the CG solve, interpolation + 8 colour sweeps -- 19 dependent launches of ~3.4 us for ~0.2 us of work each.  Here the level's
// x and r live in the shared memory of one CTA (colour-blocked layout of the level, padding included, so the offset tables of
// the per-colour kernels apply unchanged), thread t owns storage position t of every colour block and vertex t of the coarsest
// level, and the phases are separated by bar.sync: a phase is one trip to L2 for the operator row (issued with the shared-memory
// loads of x behind the barrier) plus the reference's 27 multiply-subtract pairs.  The arithmetic per vertex is the per-phase
// kernels' (same stencil order, same tables): bit-identical (tests/test_kernels_gpu.py::test_vcycle_bottom_matches_unfused).
#pragma once
#include "mg_kernels.cuh"

namespace mf {

constexpr int MG_BOTTOM_THREADS = 768;     // = k_mg_coarse_cg<.,768>: up to 85 registers per thread
struct BottomArgs {
  LvlD L;                  // the last swept level (colour-blocked, float operator, 3-D)
  LvlD B;                  // the coarsest level
  MgDims D;
  const MgOff8* off;       // device copy of L's offset tables
  float accuracy;
  int fold;                // colour 0 of the pre-smoothing was done by the restriction onto L
  int nt;                  // storage positions per colour block of L: (PX-2)(PY-2)(PZ-2) <= MG_BOTTOM_THREADS
  int vecDoubles;          // doubles of the CG search vector (with margins)
};
// dynamic shared memory, in this order: double vec[vecDoubles]; double sh[2][32]; double out[4]; float sx[L.na]; float sr[L.na];
// float sxc[B.na]; int ox[8][27]; long long oa[8][27]
__host__ __device__ inline size_t mgBottomSmem(int naL, int naB, int vecDoubles) {
  return sizeof(double) * (size_t)(vecDoubles + 64 + 4) + sizeof(float) * (size_t)(2 * naL + naB) + (sizeof(int) + sizeof(long long)) * 8 * 27 + 16;
}

template <bool WITH_CENTER>
__device__ __forceinline__ float bottomStencilSub(const char* __restrict__ ap, const float* sx, int p, const int* ox, const long long* oa, float sum) {
#pragma unroll
  for (int s = 0; s < 27; s++) {
    if (s == 13 && !WITH_CENTER) continue;
    const float a = *reinterpret_cast<const float*>(ap + oa[s]);
    sum -= a * sx[p + ox[s]];
  }
  return sum;
}

__global__ void __launch_bounds__(MG_BOTTOM_THREADS, 1) k_mg_bottom(const __grid_constant__ BottomArgs T) {
  extern __shared__ __align__(16) unsigned char smemRaw[];
  const LvlD& L = T.L; const LvlD& B = T.B;
  double* vec = reinterpret_cast<double*>(smemRaw);
  double (*sh)[32] = reinterpret_cast<double (*)[32]>(vec + T.vecDoubles);
  double* out = vec + T.vecDoubles + 64;
  float* sx = reinterpret_cast<float*>(out + 4);
  float* sr = sx + L.na;
  float* sxc = sr + L.na;
  int* ox = reinterpret_cast<int*>(sxc + B.na);
  long long* oa = reinterpret_cast<long long*>((reinterpret_cast<unsigned long long>(ox + 8 * 27) + 7ull) & ~7ull);
  const int tid = threadIdx.x, nth = blockDim.x;
  pdlTrigger();
  // the operators were last read a V-cycle ago (>= 1 GB of level-0 traffic since): pull them into L2 while the predecessor drains
  {
    const char* a = reinterpret_cast<const char*>(L.A);
    const long long bytes = 14LL * L.as * 4;
    for (long long o = (long long)tid * 128; o < bytes; o += (long long)nth * 128) asm volatile("prefetch.global.L2 [%0];" :: "l"(a + o));
    const char* ab = reinterpret_cast<const char*>(B.A);
    const long long bytesB = 14LL * B.as * 4;
    for (long long o = (long long)tid * 128; o < bytesB; o += (long long)nth * 128) asm volatile("prefetch.global.L2 [%0];" :: "l"(ab + o));
  }
  for (int i = tid; i < 8 * 27; i += nth) {
    const MgOff1& o = T.off->t[i / 27];
    ox[i] = (int)(o.xo[i % 27] >> 2); oa[i] = o.ao[i % 27];
  }
  for (int i = tid; i < B.na; i += nth) sxc[i] = 0.f;
  pdlWait();
  // x as the restriction onto this level left it (colour 0 = b / diag when folded, 0 elsewhere, padding 0)
  for (int i = tid; i < L.na; i += nth) { sx[i] = __ldcg(L.x + i); sr[i] = 0.f; }
  // storage positions of this thread's vertex in the 8 colour blocks
  const bool has = tid < T.nt;
  int p0 = 0, pstep = 0;
  if (has) {
    const int BX = L.PX - 2, BY = L.PY - 2;
    const int lx = tid % BX, u = tid / BX, ly = u % BY, lz = u / BY;
    p0 = (((lz + L.zoff) * L.NC) * L.PY + ly + 1) * L.PX + lx + 1;
    pstep = L.PY * L.PX;
  }
  __syncthreads();
  auto smooth = [&](int c) {
    if (has) {
      const int p = p0 + c * pstep;
      const char* ap = reinterpret_cast<const char*>(L.A + p);
      const float diag = *reinterpret_cast<const float*>(ap);
      const float sum = bottomStencilSub<false>(ap, sx, p, ox + c * 27, oa + c * 27, __ldcg(L.b + p));
      if (diag != 0.f) sx[p] = sum / diag;          // inactive vertices and padding have an all-zero row
    }
    __syncthreads();
  };
  for (int c = T.fold ? 1 : 0; c < 8; c++) smooth(c);
  if (has) {
    for (int c = 0; c < 8; c++) {
      const int p = p0 + c * pstep;
      sr[p] = bottomStencilSub<true>(reinterpret_cast<const char*>(L.A + p), sx, p, ox + c * 27, oa + c * 27, __ldcg(L.b + p));
    }
  }
  __syncthreads();
  // restriction onto the coarsest level (mgRestrictItem, blocked fine level): vertex tid of the coarsest level
  float bc = 0.f;
  if (tid < B.n) {
    const int vx = tid % B.sx; const int u = tid / B.sx; const int vy = u % B.sy, vz = u / B.sy;
    const float diag = B.A[mgIdx(B, vx, vy, vz)];
    float sum = 0.f;
#pragma unroll
    for (int dz = -1; dz <= 1; dz++)
#pragma unroll
      for (int dy = -1; dy <= 1; dy++)
#pragma unroll
        for (int dx = -1; dx <= 1; dx++) {
          const float rw = 1.0f / (float)(1 << ((dx & 1) + (dy & 1) + (dz & 1)));
          sum += rw * sr[mgIdx(L, vx * 2 + dx, vy * 2 + dy, vz * 2 + dz)];
        }
    if (diag != 0.f) bc = sum;                      // inactive coarse vertex: b stays 0
  }
  // coarsest level: x = 0 initial guess, b from the register, solution into sxc (and B.x)
  mgCoarseCg<false>(B, T.D, T.accuracy, sh, out, vec, true, bc, sxc);
  __syncthreads();
  // x += P x_coarse (mgInterpAddBlockedItem)
  if (has) {
    const int BX = L.PX - 2, BY = L.PY - 2;
    const int lx = tid % BX, u = tid / BX, ly = u % BY, lz = u / BY;
#pragma unroll
    for (int c = 0; c < 8; c++) {
      const int cx = c & 1, cy = (c >> 1) & 1, cz = (c >> 2) & 1;
      const int p = p0 + c * pstep;
      const bool off = L.A[p] == 0.f;
      const float xo = sx[p];
      float sum = 0.f;
#pragma unroll
      for (int dz = 0; dz <= cz; dz++)
#pragma unroll
        for (int dy = 0; dy <= cy; dy++)
#pragma unroll
          for (int dx = 0; dx <= cx; dx++) sum += sxc[mgIdx(B, lx + dx, ly + dy, lz + dz)];
      const float iw = 1.0f / (float)(1 << (cx + cy + cz));
      if (!off) sx[p] = xo + iw * sum;
    }
  }
  __syncthreads();
  for (int c = 7; c >= 0; c--) smooth(c);
  for (int i = tid; i < L.na; i += nth) L.x[i] = sx[i];
}

} // namespace mf
