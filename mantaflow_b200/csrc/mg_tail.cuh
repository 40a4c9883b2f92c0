// mg_tail.cuh -- the small levels of the V-cycle in two persistent kernels (reference: GridMg::doVCycle multigrid.cpp:448-504,
// knSmoothColor :668, knCalcResidual :739, knRestrict :904, knInterpolate :934).
//
// Below a few 10^4 vertices per colour a level's visit is pure dependency latency: 7-8 colour sweeps + residual + restriction on
// the way down, interpolation + 8 colour sweeps on the way up, each a dependent launch of ~4.4 us (programmatic dependent launch
// included) for ~0.5 us of work -- 54 launches per V-cycle at 256^3 (levels 65^3, 33^3, 17^3).  Here ONE kernel with one CTA per
// SM walks all phases of all those levels on the way down, another one on the way up (the coarsest-level CG runs between them
// as before); phases are separated by a grid barrier (one release-add on a monotonic counter + acquire spin, ~0.5 us).  The
// arithmetic per vertex is the per-phase kernels' (same stencil order, same offset tables), so results are bit-identical.
//
// Memory ordering: x / r / b of the levels are written by one CTA and read by others in the next phase.  The barrier is
// bar.sync -> (thread 0) red.release.gpu -> ld.acquire.gpu spin -> bar.sync (release / acquire are cumulative over the bar.sync), and every load of an array that is
// written inside the kernel is ld.global.cg (L2, never a stale L1 line); the operator and the offset tables are read-only.
// Co-residency: the grid is one CTA per SM with no shared memory; a CTA that is not resident yet only waits for kernels that
// finish on their own, and the dependent (PDL) launch behind this kernel cannot start before every CTA has begun.
#pragma once
#include "mg_kernels.cuh"

namespace mf {

constexpr int MG_TAIL_MAXL = 6;          // levels of one tail: MG_TAIL_MAXL - 1 swept levels + the coarsest one
constexpr int MG_TAIL_THREADS = 512;
struct TailArgs {
  int nl;                                // lv[0 .. nl-2]: swept levels (fine to coarse); lv[nl-1]: the coarsest level (CG)
  int prefetch;                          // pull the operators into L2 before the first phase
  int fold;                              // mgFoldFirst in use: colour 0 of the pre-smoothing is done by the restriction
  int cluster;                           // the grid is ONE thread-block cluster: phases are separated by the hardware cluster barrier
  unsigned base;                         // value of *bar when the kernel starts (host-tracked; all arithmetic wraps)
  unsigned* bar;                         // monotonic arrival counter
  int* err;                              // mapped host flag: a barrier timed out
  LvlD lv[MG_TAIL_MAXL];
  unsigned long long* dbg;               // development: globaltimer after every barrier (block 0), or null
  const MgOff8* off;                     // device copy of the offset tables of lv[0 .. nl-2]
};
// barriers one launch performs (host and device must agree)
__host__ __device__ inline int mgTailBarriers(int nl, bool down, bool fold) {
  const int swept = nl - 1;
  if (down) return swept * ((fold ? 7 : 8) + 2) - 1;     // per level: colour sweeps, residual, restriction; none after the last phase
  return swept * 9 - 1;                                   // per level: interpolation, 8 colour sweeps
}

__device__ __forceinline__ void tailBarrier(const TailArgs& T, unsigned& target, bool& dead) {
  if (T.cluster) {
    // release / acquire at cluster scope: the global stores of every CTA of the cluster before the barrier are visible to the
    // (L2) loads of every CTA behind it
    // (SASS: MEMBAR.ALL.GPU + UCGABAR_ARV / UCGABAR_WAIT + CCTL.IVALL.  Measured on 16 CTAs: the bare barrier 0.24 us, with
    // release / acquire 0.75 us; a smoothing phase of <= 1 vertex per thread is 1.8 us of dependent instructions on top)
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (T.dbg && blockIdx.x == 0 && threadIdx.x == 0) {
      target += gridDim.x;
      unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); T.dbg[(target - T.base) / gridDim.x] = t;
    }
    return;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    target += gridDim.x;
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(T.bar) : "memory");
    const long long t0 = clock64();
    while (!dead) {
      unsigned v;
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(T.bar) : "memory");
      if ((int)(v - target) >= 0) break;
      if (clock64() - t0 > 2000000000LL) { *T.err = 1; dead = true; }  // ~1 s: report (mg_check_errors) instead of hanging the device
    }
    if (T.dbg && blockIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); T.dbg[(target - T.base) / gridDim.x] = t; }
  }
  __syncthreads();
}

// storage position of item t among the vertices of colour c (all z pairs of the level)
__device__ __forceinline__ int tailPos(const LvlD& L, int c, int t) {
  const int BX = L.PX - 2, BY = L.PY - 2;
  const int ly0 = (BX == 1) ? t : (int)__umulhi((unsigned)t, L.bxMagic);       // t / BX
  const int lx = t - ly0 * BX;
  const int lz = ly0 / BY, ly = ly0 - lz * BY;
  return (((lz + L.zoff) * L.NC + c) * L.PY + ly + 1) * L.PX + lx + 1;
}
// The offset tables of all swept levels are staged in shared memory once (54 register-indexed constant loads per vertex would
// miss the constant cache one after the other; 54 broadcast LDS do not).  A phase is latency: every round trip to L2 costs
// ~0.35 us on this part, a barrier ~1.25 us (tools/barrier_probe.cu), so the items issue ALL their loads (diagonal, b, 26-27
// coefficients, 26-27 x) in one batch and predicate the store instead of testing the row first.
template <bool WITH_CENTER>
__device__ __forceinline__ float tailStencilSub(const char* __restrict__ ap, const char* xp, const long long* o, float sum) {
#pragma unroll
  for (int s = 0; s < 27; s++) {
    if (s == 13 && !WITH_CENTER) continue;
    const float a = *reinterpret_cast<const float*>(ap + o[27 + s]);
    sum -= a * __ldcg(reinterpret_cast<const float*>(xp + o[s]));
  }
  return sum;
}
__device__ __forceinline__ void tailSmooth(const LvlD& L, const long long* o, int c, int tid, int nth) {
  const int nt = (L.PX - 2) * (L.PY - 2) * (L.PZ - 2);
  for (int t = tid; t < nt; t += nth) {
    const int p = tailPos(L, c, t);
    const char* ap = reinterpret_cast<const char*>(L.A + p);
    const float diag = *reinterpret_cast<const float*>(ap);
    const float sum = tailStencilSub<false>(ap, reinterpret_cast<const char*>(L.x + p), o, __ldcg(L.b + p));
    if (diag != 0.f) L.x[p] = sum / diag;          // inactive vertices and padding have an all-zero row
  }
}
// all 8 colours at once: item = (colour, position)
__device__ __forceinline__ void tailResidual(const LvlD& L, const long long* o8, int tid, int nth) {
  const int nt = (L.PX - 2) * (L.PY - 2) * (L.PZ - 2);
  for (int i = tid; i < 8 * nt; i += nth) {
    const int c = i / nt, t = i - c * nt;
    const int p = tailPos(L, c, t);
    L.r[p] = tailStencilSub<true>(reinterpret_cast<const char*>(L.A + p), reinterpret_cast<const char*>(L.x + p), o8 + 54 * c, __ldcg(L.b + p));
  }
}
// mgRestrictItem<true> (blocked fine level) with L2 loads of the fine residual
__device__ __forceinline__ void tailRestrict(const LvlD& F, const LvlD& C, bool fold, int tid, int nth) {
  for (int t = tid; t < C.n; t += nth) {
    const int vx = t % C.sx; const int u = t / C.sx; const int vy = u % C.sy, vz = u / C.sy;
    const int v = mgIdx(C, vx, vy, vz);
    const float diag = C.A[v];
    float sum = 0.f;
#pragma unroll
    for (int dz = -1; dz <= 1; dz++)
#pragma unroll
      for (int dy = -1; dy <= 1; dy++)
#pragma unroll
        for (int dx = -1; dx <= 1; dx++) {
          const float rw = 1.0f / (float)(1 << ((dx & 1) + (dy & 1) + (dz & 1)));
          sum += rw * __ldcg(F.r + mgIdx(F, vx * 2 + dx, vy * 2 + dy, vz * 2 + dz));
        }
    if (diag == 0.f) { C.x[v] = 0.f; continue; }          // inactive coarse vertex: b stays 0
    C.b[v] = sum;
    C.x[v] = (fold && ((vx | vy | vz) & 1) == 0) ? sum / diag : 0.f;
  }
}
// mgInterpAddItem with L2 loads
__device__ __forceinline__ void tailInterpAdd(const LvlD& F, const LvlD& C, int tid, int nth) {
  for (int t = tid; t < F.n; t += nth) {
    const int vx = t % F.sx; const int u = t / F.sx; const int vy = u % F.sy, vz = u / F.sy;
    const int v = mgIdx(F, vx, vy, vz);
    const float diag = F.A[v];
    const float xo = __ldcg(F.x + v);
    float sum = 0.f;
    for (int z = vz / 2; z <= (vz + 1) / 2; z++)
      for (int y = vy / 2; y <= (vy + 1) / 2; y++)
        for (int x = vx / 2; x <= (vx + 1) / 2; x++) sum += __ldcg(C.x + mgIdx(C, x, y, z));
    const float iw = 1.0f / (float)(1 << ((vx % 2) + (vy % 2) + (vz % 2)));
    if (diag != 0.f) F.x[v] = xo + iw * sum;
  }
}

template <bool DOWN>
__global__ void __launch_bounds__(MG_TAIL_THREADS, 1) k_mg_tail(const __grid_constant__ TailArgs T) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  pdlTrigger();
  // the operators of the swept levels are constant and were last read a V-cycle ago (>= 1 GB of level-0 traffic since): pull
  // them into L2 while the predecessor kernel drains, so the phases do not start with a DRAM miss each
  if (T.prefetch) {
    for (int li = 0; li < T.nl - 1; li++) {
      const char* a = reinterpret_cast<const char*>(T.lv[li].A);
      const long long bytes = 14LL * T.lv[li].as * 4;
      for (long long o = (long long)tid * 128; o < bytes; o += (long long)nth * 128) asm volatile("prefetch.global.L2 [%0];" :: "l"(a + o));
    }
  }
  pdlWait();
  __shared__ long long sOff[(MG_TAIL_MAXL - 1) * 8 * 54];
  unsigned target = T.base;
  bool dead = false;
  if (T.dbg && blockIdx.x == 0 && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); T.dbg[0] = t; }
  const int swept = T.nl - 1;
  {
    const long long* src = reinterpret_cast<const long long*>(T.off);       // MgOff1 = xo[27], ao[27]
    for (int i = threadIdx.x; i < swept * 8 * 54; i += blockDim.x) sOff[i] = src[i];
    __syncthreads();
  }
  if (DOWN) {
    for (int li = 0; li < swept; li++) {
      const LvlD& L = T.lv[li];
      for (int c = T.fold ? 1 : 0; c < 8; c++) { tailSmooth(L, sOff + (li * 8 + c) * 54, c, tid, nth); tailBarrier(T, target, dead); }
      tailResidual(L, sOff + li * 8 * 54, tid, nth);
      tailBarrier(T, target, dead);
      tailRestrict(L, T.lv[li + 1], T.fold && li + 1 < swept, tid, nth);
      if (li + 1 < swept) tailBarrier(T, target, dead);
    }
  } else {
    for (int li = swept - 1; li >= 0; li--) {
      const LvlD& L = T.lv[li];
      tailInterpAdd(L, T.lv[li + 1], tid, nth);
      tailBarrier(T, target, dead);
      for (int c = 7; c >= 0; c--) { tailSmooth(L, sOff + (li * 8 + c) * 54, c, tid, nth); if (c > 0 || li > 0) tailBarrier(T, target, dead); }
    }
  }
}

} // namespace mf
