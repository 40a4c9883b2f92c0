// mic.cuh -- the reference's sequential modified incomplete Cholesky preconditioner MIC(0), bit for bit, as a tile wavefront.
// Reference: InitPreconditionModifiedIncompCholesky2 (source/conjugategrad.cpp:66-96) and
// ApplyPreconditionModifiedIncompCholesky2 (:135-158): three lexicographic sweeps (set-up, forward substitution, backward
// substitution) in which cell (i,j,k) needs the results of (i-1,j,k), (i,j-1,k), (i,j,k-1) (backward: the +1 neighbours).
//
// The sweeps keep the reference's arithmetic expression by expression (float products in its association order, the one
// double-precision step of the set-up, 1.0f / sqrtf), so PcMIC takes the reference's iterates and iteration counts; only the
// ORDER in which independent cells are visited differs: the grid is cut into 16 x 16 x TZ tiles, a tile is swept by one CTA along
// its hyperplanes i+j+k = const out of shared memory (one thread per x-line, one block barrier per hyperplane), and tiles run as a
// wavefront of their own: persistent CTAs draw tiles from a ticket counter in wavefront order (bi+bj+bk ascending) and wait on the
// completion flags of their three predecessor tiles.  A CTA only ever waits for tiles with smaller tickets, which are held by CTAs
// that are already running, so the scheme cannot deadlock whatever the grid size.
#pragma once
#include "common.cuh"
#include "cg.cuh"

namespace mf {

struct MicPlan {
  int tx, ty, tz, ntiles, tzEdge;   // tiles per axis; tile depth
  int* order;     // device: tile ids in wavefront order
  int* done;      // device: per tile, the epoch of the sweep that last finished it
  int* ctl;       // device: [0] ticket counter, [1] finished CTAs (the last one resets both)
  int epoch;
  int ctas;
  MicPlan* code16;  // the 16^3-tile plan of the row-code fast path (k_mic_code), lazily created
};

struct MicArgs {
  const int* flags; const float *A0, *Ai, *Aj, *Ak; float* P; const float* src; float* dst;
  const int* order; int* done; int* ctl; int ntiles, tx, ty, tz, epoch; const CgState* st;
};

__device__ __forceinline__ int micLdAcquire(const int* p) { int v; asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void micStRelease(int* p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }

// MODE 0: set-up (writes P = Aprecond), 1: forward substitution dst = L^-1 src, 2: backward substitution in place on dst
template <int MODE, int TZ>
__global__ void __launch_bounds__(16 * TZ) k_mic_wave(Geo g, MicArgs a) {
  constexpr int TX = 16, TY = 16, S = TX + 3, R = TY + 2, PL = S * R, NT = TY * TZ, SZ = PL * (TZ + 1);
  extern __shared__ float sm[];
  float* sAi = sm; float* sAj = sm + SZ; float* sAk = sm + 2 * SZ; float* sP = sm + 3 * SZ; float* sD = sm + 4 * SZ;
  __shared__ int sQ;
  const int tid = threadIdx.x;
  if (MODE != 0 && a.st->done) return;
  const int jl = tid % TY, kl = tid / TY;
  const int dir = (MODE == 2) ? -1 : 1;
  const long long Y = g.nx, Z = g.Z;
  for (;;) {
    if (tid == 0) sQ = atomicAdd(&a.ctl[0], 1);
    __syncthreads();
    const int q = sQ;
    if (q >= a.ntiles) break;
    int t = a.order[q];
    int bi = t % a.tx, bj = (t / a.tx) % a.ty, bk = t / (a.tx * a.ty);
    if (MODE == 2) { bi = a.tx - 1 - bi; bj = a.ty - 1 - bj; bk = a.tz - 1 - bk; t = bi + a.tx * (bj + a.ty * bk); }
    // local (il, jl, kl) -> global (i0 + dir * il, ...): local order is the sweep order
    const int i0 = (MODE == 2) ? bi * TX + TX - 1 : bi * TX;
    const int j0 = (MODE == 2) ? bj * TY + TY - 1 : bj * TY;
    const int k0 = (MODE == 2) ? bk * TZ + TZ - 1 : bk * TZ;
    // ---- static data of the tile and of its lower halo (nothing here depends on other tiles of this sweep)
    for (int c = tid; c < (TX + 1) * (TY + 1) * (TZ + 1); c += NT) {
      const int li = c % (TX + 1) - 1, lj = (c / (TX + 1)) % (TY + 1) - 1, lk = c / ((TX + 1) * (TY + 1)) - 1;
      const int gi = i0 + dir * li, gj = j0 + dir * lj, gk = k0 + dir * lk;
      const bool in = gi >= 0 && gi < g.nx && gj >= 0 && gj < g.ny && gk >= 0 && gk < g.nz;
      const long long idx = gi + Y * gj + Z * gk;
      const int sc = (li + 1) + S * ((lj + 1) + R * (lk + 1));
      sAi[sc] = in ? a.Ai[idx] : 0.f; sAj[sc] = in ? a.Aj[idx] : 0.f; sAk[sc] = in ? a.Ak[idx] : 0.f;
      const bool interior = li >= 0 && lj >= 0 && lk >= 0;
      if (MODE == 0) { sD[sc] = in ? a.A0[idx] : 0.f; sP[sc] = 0.f; }
      else {
        sP[sc] = in ? a.P[idx] : 0.f;
        if (MODE == 1) sD[sc] = (interior && in && fIsFluid(a.flags[idx])) ? a.src[idx] : 0.f;   // non-fluid cells: the cleared temp grid's 0
        else sD[sc] = (interior && in) ? a.dst[idx] : 0.f;
      }
    }
    // fluid cells of this thread's x-line
    unsigned mask = 0;
    {
      const int gj = j0 + dir * jl, gk = k0 + dir * kl;
      if (gj >= 0 && gj < g.ny && gk >= 0 && gk < g.nz) {
        for (int il = 0; il < TX; il++) { const int gi = i0 + dir * il; if (gi >= 0 && gi < g.nx && fIsFluid(a.flags[gi + Y * gj + Z * gk])) mask |= 1u << il; }
      }
    }
    // ---- wait for the three predecessor tiles, then fetch the faces they produced (L2 loads: another SM wrote them)
    if (tid < 3) {
      const int pb = (tid == 0) ? bi - dir : (tid == 1) ? bj - dir : bk - dir;
      const int lim = (tid == 0) ? a.tx : (tid == 1) ? a.ty : a.tz;
      if (pb >= 0 && pb < lim) {
        const int pt = (tid == 0) ? t - dir : (tid == 1) ? t - dir * a.tx : t - dir * a.tx * a.ty;
        while (micLdAcquire(&a.done[pt]) != a.epoch) { }
      }
    }
    __syncthreads();
    {
      float* sdyn = (MODE == 0) ? sP : sD;
      const float* gdyn = (MODE == 0) ? a.P : a.dst;
      for (int c = tid; c < TY * TZ + TX * TZ + TX * TY; c += NT) {
        int li, lj, lk;
        if (c < TY * TZ) { li = -1; lj = c % TY; lk = c / TY; }
        else if (c < TY * TZ + TX * TZ) { const int d = c - TY * TZ; li = d % TX; lj = -1; lk = d / TX; }
        else { const int d = c - TY * TZ - TX * TZ; li = d % TX; lj = d / TX; lk = -1; }
        const int gi = i0 + dir * li, gj = j0 + dir * lj, gk = k0 + dir * lk;
        const bool in = gi >= 0 && gi < g.nx && gj >= 0 && gj < g.ny && gk >= 0 && gk < g.nz;
        sdyn[(li + 1) + S * ((lj + 1) + R * (lk + 1))] = in ? __ldcg(gdyn + gi + Y * gj + Z * gk) : 0.f;
      }
    }
    __syncthreads();
    // ---- hyperplane sweep of the tile
    for (int h = 0; h < TX + TY + TZ - 2; h++) {
      const int il = h - jl - kl;
      if (il >= 0 && il < TX && ((mask >> il) & 1u)) {
        const int c = (il + 1) + S * ((jl + 1) + R * (kl + 1));
        if (MODE == 0) {
          // conjugategrad.cpp:79-95; `tau * ( ... + 0.)` is evaluated in double
          const float px = sP[c - 1], py = sP[c - S], pz = sP[c - PL];
          const float ax = sAi[c - 1], ay = sAj[c - S], az = sAk[c - PL];
          const float a0 = sD[c];
          const float qx = ax * px, qy = ay * py, qz = az * pz;
          float e = a0 - qx * qx;
          e = e - qy * qy;
          e = e - qz * qz;
          float sum = (ax * (sAj[c - 1] + sAk[c - 1])) * (px * px);
          sum = sum + (ay * (sAi[c - S] + sAk[c - S])) * (py * py);
          sum = sum + (az * (sAi[c - PL] + sAj[c - PL])) * (pz * pz);
          e = (float)((double)e - (double)0.97f * ((double)sum + 0.));
          if (e < 0.25f * a0) e = a0;
          sP[c] = 1.0f / sqrtf(e);
        } else if (MODE == 1) {
          // dst = p * (src - dst(i-1) * Ai(i-1) * P(i-1) - ...)      (conjugategrad.cpp:141-148)
          const float p = sP[c];
          float v = sD[c];
          v = v - (sD[c - 1] * sAi[c - 1]) * sP[c - 1];
          v = v - (sD[c - S] * sAj[c - S]) * sP[c - S];
          v = v - (sD[c - PL] * sAk[c - PL]) * sP[c - PL];
          sD[c] = p * v;
        } else {
          // dst = p * (dst - dst(i+1) * Ai * p - ...)                 (conjugategrad.cpp:151-158)
          const float p = sP[c];
          float v = sD[c];
          v = v - (sD[c - 1] * sAi[c]) * p;
          v = v - (sD[c - S] * sAj[c]) * p;
          v = v - (sD[c - PL] * sAk[c]) * p;
          sD[c] = p * v;
        }
      }
      __syncthreads();
    }
    // ---- results to global memory, then publish the tile
    {
      const float* sres = (MODE == 0) ? sP : sD;
      float* gres = (MODE == 0) ? a.P : a.dst;
      for (int c = tid; c < TX * TY * TZ; c += NT) {
        const int li = c % TX, lj = (c / TX) % TY, lk = c / (TX * TY);
        const int gi = i0 + dir * li, gj = j0 + dir * lj, gk = k0 + dir * lk;
        if (gi >= 0 && gi < g.nx && gj >= 0 && gj < g.ny && gk >= 0 && gk < g.nz) gres[gi + Y * gj + Z * gk] = sres[(li + 1) + S * ((lj + 1) + R * (lk + 1))];
      }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) micStRelease(&a.done[t], a.epoch);
  }
  if (tid == 0) { if (atomicAdd(&a.ctl[1], 1) == (int)gridDim.x - 1) { a.ctl[0] = 0; a.ctl[1] = 0; __threadfence(); } }
}

// ---- fast path for the plain Laplace matrix (16-bit row codes, pressure_kernels.cuh CGC_*) ---------------------------------
// Couplings are 0 or -1, so (dst(n) * A) * P == dst(n) * (A * P) bit for bit (a multiplication by -1 or +0 is exact), and
// A * P is -P(n) where the code has the coupling bit and +0 elsewhere: a tile needs P, the right-hand side and 2 bytes of code per
// cell (14 B/cell of DRAM traffic per sweep instead of 28).  One thread per x-line as above, but the value of the line's
// previous cell stays in a register, the coefficients of the next cell are fetched while the thread waits at the barrier, the
// results leave for global memory as they are produced (only the fence is left at the end of the tile), the ticket of the next
// tile is drawn while this one is swept, and the three upstream faces are fetched by three warps that each wait for their own flag.
// MODE 1 forward, 2 backward.  Tile 16 x 16 x 16, 256 threads.
struct MicCodeArgs {
  const unsigned short* code; const float* P; const float* src; float* dst;
  const int* order; int* done; int* ctl; int ntiles, tx, ty, tz, epoch; const CgState* st;
  unsigned char lines[256];   // thread -> x-line (jl | kl << 4), sorted by jl + kl
  unsigned long long* dbg;    // MF_MIC_DBG: 8 globaltimer stamps per tile (ticket, static data, flags seen, faces, sweep, published, SM id)
};
__device__ __forceinline__ unsigned long long micNow() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define MIC_STAMP(slot) do { if (a.dbg && tid == 0) a.dbg[(size_t)t * 8 + (slot)] = micNow(); } while (0)
template <int MODE>
__global__ void __launch_bounds__(256, 3) k_mic_code(Geo g, MicCodeArgs a) {
  constexpr int TX = 16, TY = 16, TZ = 16, S = TX + 3, R = TY + 2, PL = S * R, SZ = PL * (TZ + 1);
  extern __shared__ float sm[];
  float* sP = sm; float* sD = sm + SZ;
  unsigned short* sC = reinterpret_cast<unsigned short*>(sm + 2 * SZ);
  __shared__ int sQ[2];
  const int tid = threadIdx.x;
  if (a.st->done) return;
  // lines in the order of their diagonal jl + kl: the lines that are active on a hyperplane sit in neighbouring warps, the
  // other warps only pass the barrier
  const int jl = a.lines[tid] & 15, kl = a.lines[tid] >> 4;
  constexpr int dir = (MODE == 2) ? -1 : 1;
  constexpr int BI = (MODE == 2) ? CGC_XP : CGC_XM, BJ = (MODE == 2) ? CGC_YP : CGC_YM, BK = (MODE == 2) ? CGC_ZP : CGC_ZM;
  const long long Y = g.nx, Z = g.Z;
  const float* gsrc = (MODE == 1) ? a.src : a.dst;
  if (tid == 0) sQ[0] = atomicAdd(&a.ctl[0], 1);
  __syncthreads();
  for (int round = 0;; round++) {
    const int q = sQ[round & 1];
    if (q >= a.ntiles) break;
    int nextTicket = 0;
    if (tid == 0) nextTicket = atomicAdd(&a.ctl[0], 1);     // consumed at the end of this tile
    int t = a.order[q];
    int bi = t % a.tx, bj = (t / a.tx) % a.ty, bk = t / (a.tx * a.ty);
    if (MODE == 2) { bi = a.tx - 1 - bi; bj = a.ty - 1 - bj; bk = a.tz - 1 - bk; t = bi + a.tx * (bj + a.ty * bk); }
    const int i0 = (MODE == 2) ? bi * TX + TX - 1 : bi * TX;      // local (il, jl, kl) -> global (i0 + dir * il, ...)
    const int j0 = (MODE == 2) ? bj * TY + TY - 1 : bj * TY;
    const int k0 = (MODE == 2) ? bk * TZ + TZ - 1 : bk * TZ;
    MIC_STAMP(0);
    if (a.dbg && tid == 0) { unsigned smid; asm volatile("mov.u32 %0, %smid;" : "=r"(smid)); a.dbg[(size_t)t * 8 + 6] = smid; a.dbg[(size_t)t * 8 + 7] = q; }
    // ---- static data, all loads of a thread in one batch: the tile's rows are 16 floats = 4 x 16 bytes (nx % 4 == 0)
    {
      const int xb = bi * TX;
      float4 vp[4], vs[4]; uint2 vc[4]; float hp[3];
      int rowOk[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {           // (row, quarter): 1024 items
        const int item = tid + 256 * u, qx = item & 3, row = item >> 2, rj = row & 15, rk = row >> 4;     // row in GLOBAL orientation
        const int gj = bj * TY + rj, gk = bk * TZ + rk, gx = xb + 4 * qx;
        rowOk[u] = (gj < g.ny && gk < g.nz && gx < g.nx);
        const long long idx = gx + Y * gj + Z * gk;
        vp[u] = make_float4(0.f, 0.f, 0.f, 0.f); vs[u] = vp[u]; vc[u] = make_uint2(0u, 0u);
        if (rowOk[u]) {
          vp[u] = *reinterpret_cast<const float4*>(a.P + idx);
          vs[u] = *reinterpret_cast<const float4*>(gsrc + idx);
          vc[u] = *reinterpret_cast<const uint2*>(a.code + idx);
        }
      }
      if (MODE == 1) {                         // P on the three lower faces (768 values)
#pragma unroll
        for (int u = 0; u < 3; u++) {
          const int c = tid;
          int li, lj, lk;
          if (u == 0) { li = -1; lj = c % TY; lk = c / TY; }
          else if (u == 1) { li = c % TX; lj = -1; lk = c / TX; }
          else { li = c % TX; lj = c / TX; lk = -1; }
          const int gi = i0 + li, gj = j0 + lj, gk = k0 + lk;
          const bool in = gi >= 0 && gi < g.nx && gj >= 0 && gj < g.ny && gk >= 0 && gk < g.nz;
          hp[u] = in ? a.P[gi + Y * gj + Z * gk] : 0.f;
        }
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int item = tid + 256 * u, qx = item & 3, row = item >> 2, rj = row & 15, rk = row >> 4;
        const int lj = (MODE == 2) ? TY - 1 - rj : rj, lk = (MODE == 2) ? TZ - 1 - rk : rk;
        const float pv[4] = { vp[u].x, vp[u].y, vp[u].z, vp[u].w };
        const float sv[4] = { vs[u].x, vs[u].y, vs[u].z, vs[u].w };
        const unsigned cv[4] = { vc[u].x & 0xffffu, vc[u].x >> 16, vc[u].y & 0xffffu, vc[u].y >> 16 };
#pragma unroll
        for (int e = 0; e < 4; e++) {
          const int xi = 4 * qx + e, li = (MODE == 2) ? TX - 1 - xi : xi;
          const int sc = (li + 1) + S * ((lj + 1) + R * (lk + 1));
          sP[sc] = pv[e];
          sD[sc] = (cv[e] & CGC_FLUID) ? sv[e] : 0.f;       // non-fluid cells: the cleared temp grid's 0
          sC[li + TX * (lj + TY * lk)] = (unsigned short)cv[e];
        }
      }
      // halo slots: P of the lower faces (forward), zero for the dynamic values until the upstream faces arrive
#pragma unroll
      for (int u = 0; u < 3; u++) {
        const int c = tid;
        int li, lj, lk;
        if (u == 0) { li = -1; lj = c % TY; lk = c / TY; }
        else if (u == 1) { li = c % TX; lj = -1; lk = c / TX; }
        else { li = c % TX; lj = c / TX; lk = -1; }
        const int sc = (li + 1) + S * ((lj + 1) + R * (lk + 1));
        if (MODE == 1) sP[sc] = hp[u];
        sD[sc] = 0.f;
      }
    }
    __syncthreads();
    MIC_STAMP(1);
    // ---- the three upstream faces, each fetched by the warp that waited for its tile (8 values per lane, one batch)
    {
      const int w = tid >> 5, lane = tid & 31;
      if (w < 3) {
        const int pb = (w == 0) ? bi - dir : (w == 1) ? bj - dir : bk - dir;
        const int lim = (w == 0) ? a.tx : (w == 1) ? a.ty : a.tz;
        if (pb >= 0 && pb < lim) {
          const int pt = (w == 0) ? t - dir : (w == 1) ? t - dir * a.tx : t - dir * a.tx * a.ty;
          if (lane == 0) { while (micLdAcquire(&a.done[pt]) != a.epoch) { } }
          __syncwarp();
          if (a.dbg && lane == 0) atomicMax(&a.dbg[(size_t)t * 8 + 2], micNow());
          float fv[8]; int fs[8];
#pragma unroll
          for (int u = 0; u < 8; u++) {
            const int c = lane + 32 * u;
            int li, lj, lk;
            if (w == 0) { li = -1; lj = c % TY; lk = c / TY; }
            else if (w == 1) { li = c % TX; lj = -1; lk = c / TX; }
            else { li = c % TX; lj = c / TX; lk = -1; }
            const int gi = i0 + dir * li, gj = j0 + dir * lj, gk = k0 + dir * lk;
            const bool in = gi >= 0 && gi < g.nx && gj >= 0 && gj < g.ny && gk >= 0 && gk < g.nz;
            fs[u] = (li + 1) + S * ((lj + 1) + R * (lk + 1));
            fv[u] = in ? __ldcg(a.dst + gi + Y * gj + Z * gk) : 0.f;
          }
#pragma unroll
          for (int u = 0; u < 8; u++) sD[fs[u]] = fv[u];
        }
      }
    }
    __syncthreads();
    MIC_STAMP(3);
    // ---- hyperplane sweep: this thread's line is (jl, kl), its cell on hyperplane h is il = h - jl - kl
    {
      const int cb = 1 + S * ((jl + 1) + R * (kl + 1));          // smem index of il = 0
      const unsigned short* cl = sC + TX * (jl + TY * kl);
      const int hs = jl + kl;
      // warp-uniform range of hyperplanes on which some lane of this warp has a cell
      const int wlo = __reduce_min_sync(0xffffffffu, hs), whi = __reduce_max_sync(0xffffffffu, hs) + TX - 1;
      float dprev = sD[cb - 1];
      int cd = cl[0];
      float p = sP[cb], v0 = sD[cb];
      float ci = (cd & BI) ? -((MODE == 1) ? sP[cb - 1] : p) : 0.f;
      float cj = (cd & BJ) ? -((MODE == 1) ? sP[cb - S] : p) : 0.f;
      float ck = (cd & BK) ? -((MODE == 1) ? sP[cb - PL] : p) : 0.f;
#pragma unroll 2
      for (int h = 0; h < TX + TY + TZ - 2; h++) {
        if (h >= wlo && h <= whi) {
          const int il = h - hs;
          if (il >= 0 && il < TX) {
            const int c = cb + il;
            const int cn = (il + 1 < TX) ? c + 1 : c;            // next cell of the line (clamped: the values are then unused)
            // all loads first: the two neighbours this step depends on, then the next cell's static data
            const float dj = sD[c - S], dk = sD[c - PL];
            const int cdn = cl[(il + 1 < TX) ? il + 1 : il];
            const float pn = sP[cn], v0n = sD[cn];
            const float pjn = (MODE == 1) ? sP[cn - S] : 0.f, pkn = (MODE == 1) ? sP[cn - PL] : 0.f;
            float v = v0 - dprev * ci;
            v = v - dj * cj;
            v = v - dk * ck;
            const float d = (cd & CGC_FLUID) ? p * v : 0.f;
            sD[c] = d; dprev = d;
            ci = (cdn & BI) ? -((MODE == 1) ? p : pn) : 0.f;
            cj = (cdn & BJ) ? -((MODE == 1) ? pjn : pn) : 0.f;
            ck = (cdn & BK) ? -((MODE == 1) ? pkn : pn) : 0.f;
            p = pn; v0 = v0n; cd = cdn;
          }
        }
        __syncthreads();
      }
    }
    MIC_STAMP(4);
    // ---- results to global memory in whole 16-byte pieces (scattered 4-byte stores from inside the sweep would queue in front of
    // its shared-memory loads), then publish the tile
    {
      const int xb = bi * TX;
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int item = tid + 256 * u, qx = item & 3, row = item >> 2, rj = row & 15, rk = row >> 4;
        const int gj = bj * TY + rj, gk = bk * TZ + rk, gx = xb + 4 * qx;
        if (gj < g.ny && gk < g.nz && gx < g.nx) {
          const int lj = (MODE == 2) ? TY - 1 - rj : rj, lk = (MODE == 2) ? TZ - 1 - rk : rk;
          const int rb = 1 + S * ((lj + 1) + R * (lk + 1));
          float4 o;
          if (MODE == 2) { const int l0 = TX - 1 - 4 * qx; o = make_float4(sD[rb + l0], sD[rb + l0 - 1], sD[rb + l0 - 2], sD[rb + l0 - 3]); }
          else { const int l0 = 4 * qx; o = make_float4(sD[rb + l0], sD[rb + l0 + 1], sD[rb + l0 + 2], sD[rb + l0 + 3]); }
          *reinterpret_cast<float4*>(a.dst + gx + Y * gj + Z * gk) = o;
        }
      }
    }
    // every thread fences its own stores: measured on B200, leaving the flush to thread 0's cumulative release alone costs 23 us
    // per tile instead of 1.7 us (gpurun_out/mic_dbg.bin, tools/mic_dbg.py)
    __threadfence();
    if (tid == 0) sQ[(round + 1) & 1] = nextTicket;
    __syncthreads();
    if (tid == 0) micStRelease(&a.done[t], a.epoch);
    MIC_STAMP(5);
  }
  if (tid == 0) { if (atomicAdd(&a.ctl[1], 1) == (int)gridDim.x - 1) { a.ctl[0] = 0; a.ctl[1] = 0; __threadfence(); } }
}
inline size_t micCodeSmemBytes() { return sizeof(float) * 2 * (size_t)(19 * 18 * 17) + sizeof(unsigned short) * 16 * 16 * 16; }

template <int TZ> inline size_t micSmemBytes() { return sizeof(float) * 5 * (size_t)(19 * 18 * (TZ + 1)); }

inline void micDestroy(MicPlan* m) {
  if (!m) return;
  if (m->code16) micDestroy(m->code16);
  cudaFree(m->order); cudaFree(m->done); cudaFree(m->ctl);
  delete m;
}

inline MicPlan* micBuild(Solver* s, int tzEdge) {
  MicPlan* m = new MicPlan();
  m->code16 = nullptr;
  m->tzEdge = tzEdge;
  m->tx = (s->nx + 15) / 16; m->ty = (s->ny + 15) / 16; m->tz = (s->nz + m->tzEdge - 1) / m->tzEdge;
  m->ntiles = m->tx * m->ty * m->tz;
  std::vector<int> order(m->ntiles);
  {   // counting sort by wavefront index bi + bj + bk
    const int nw = m->tx + m->ty + m->tz - 2;
    std::vector<int> start(nw + 1, 0);
    for (int t = 0; t < m->ntiles; t++) { const int w = t % m->tx + (t / m->tx) % m->ty + t / (m->tx * m->ty); start[w + 1]++; }
    for (int w = 0; w < nw; w++) start[w + 1] += start[w];
    for (int t = 0; t < m->ntiles; t++) { const int w = t % m->tx + (t / m->tx) % m->ty + t / (m->tx * m->ty); order[start[w]++] = t; }
  }
  CUDA_CHECK(cudaMalloc(&m->order, sizeof(int) * m->ntiles));
  CUDA_CHECK(cudaMalloc(&m->done, sizeof(int) * m->ntiles));
  CUDA_CHECK(cudaMalloc(&m->ctl, sizeof(int) * 2));
  CUDA_CHECK(cudaMemcpyAsync(m->order, order.data(), sizeof(int) * m->ntiles, cudaMemcpyHostToDevice, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));   // `order` leaves scope
  CUDA_CHECK(cudaMemsetAsync(m->done, 0, sizeof(int) * m->ntiles, s->stream));
  CUDA_CHECK(cudaMemsetAsync(m->ctl, 0, sizeof(int) * 2, s->stream));
  m->epoch = 0;
  m->ctas = 1;
  return m;
}

inline MicPlan* micPlan(Solver* s) {
  if (s->mic) return s->mic;
  MicPlan* m = micBuild(s, 8);
  int occ = 1;
  CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<0, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<8>()));
  CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<1, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<8>()));
  CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<2, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<8>()));
  CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_mic_wave<1, 8>, 128, micSmemBytes<8>()));
  if (occ < 1) occ = 1;
  m->ctas = s->numSMs * occ; if (m->ctas > m->ntiles) m->ctas = m->ntiles;
  s->mic = m;
  return m;
}

inline MicPlan* micPlanCode(Solver* s) {
  MicPlan* base = micPlan(s);
  if (base->code16) return base->code16;
  MicPlan* m = micBuild(s, 16);
  int occ = 1;
  CUDA_CHECK(cudaFuncSetAttribute(k_mic_code<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micCodeSmemBytes()));
  CUDA_CHECK(cudaFuncSetAttribute(k_mic_code<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micCodeSmemBytes()));
  CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_mic_code<1>, 256, micCodeSmemBytes()));
  if (occ < 1) occ = 1;
  if (getenv("MF_MIC_OCC")) { const int o = atoi(getenv("MF_MIC_OCC")); if (o >= 1 && o < occ) occ = o; }
  m->ctas = s->numSMs * occ; if (m->ctas > m->ntiles) m->ctas = m->ntiles;
  base->code16 = m;
  return m;
}

// forward (mode 1: dst = L^-1 src) / backward (mode 2, in place) sweep from the 16-bit row codes
inline void micSweepCode(Solver* s, int mode, const unsigned short* code, const float* P, const float* src, float* dst, const CgState* st) {
  MicPlan* m = micPlanCode(s);
  MicCodeArgs a; a.code = code; a.P = P; a.src = src; a.dst = dst;
  a.order = m->order; a.done = m->done; a.ctl = m->ctl; a.ntiles = m->ntiles; a.tx = m->tx; a.ty = m->ty; a.tz = m->tz; a.st = st;
  a.epoch = ++m->epoch;
  { int n = 0; for (int d = 0; d < 31; d++) for (int jl = 0; jl < 16; jl++) { const int kl = d - jl; if (kl >= 0 && kl < 16) a.lines[n++] = (unsigned char)(jl | (kl << 4)); } }
  a.dbg = nullptr;
  if (getenv("MF_MIC_DBG")) {     // development aid: dump the stamps of sweep number MF_MIC_DBG to mic_dbg.bin
    static unsigned long long* dbgBuf = nullptr; static int calls = 0;
    if (!dbgBuf) CUDA_CHECK(cudaMalloc(&dbgBuf, sizeof(unsigned long long) * 8 * (size_t)m->ntiles));
    if (++calls == atoi(getenv("MF_MIC_DBG"))) {
      CUDA_CHECK(cudaMemsetAsync(dbgBuf, 0, sizeof(unsigned long long) * 8 * (size_t)m->ntiles, s->stream));
      a.dbg = dbgBuf;
      const Geo g2 = geo(s);
      if (mode == 1) k_mic_code<1><<<m->ctas, 256, micCodeSmemBytes(), s->stream>>>(g2, a);
      else k_mic_code<2><<<m->ctas, 256, micCodeSmemBytes(), s->stream>>>(g2, a);
      std::vector<unsigned long long> h(8 * (size_t)m->ntiles);
      CUDA_CHECK(cudaMemcpyAsync(h.data(), dbgBuf, h.size() * 8, cudaMemcpyDeviceToHost, s->stream));
      CUDA_CHECK(cudaStreamSynchronize(s->stream));
      FILE* f = fopen("mic_dbg.bin", "wb");
      if (f) { int hdr[4] = { m->tx, m->ty, m->tz, mode }; fwrite(hdr, sizeof(int), 4, f); fwrite(h.data(), 8, h.size(), f); fclose(f); }
      s->launches++;
      return;
    }
  }
  const Geo g = geo(s);
  ProfScope ps(s, mode == 1 ? "k_mic_code_forward" : "k_mic_code_backward");
  if (mode == 1) k_mic_code<1><<<m->ctas, 256, micCodeSmemBytes(), s->stream>>>(g, a);
  else k_mic_code<2><<<m->ctas, 256, micCodeSmemBytes(), s->stream>>>(g, a);
  s->launches++;
}

// one sweep: mode 0 set-up, 1 forward (dst = L^-1 src), 2 backward (in place)
inline void micSweep(Solver* s, int mode, const int* flags, const float* A0, const float* Ai, const float* Aj, const float* Ak, float* P,
                     const float* src, float* dst, const CgState* st) {
  MicPlan* m = micPlan(s);
  MicArgs a; a.flags = flags; a.A0 = A0; a.Ai = Ai; a.Aj = Aj; a.Ak = Ak; a.P = P; a.src = src; a.dst = dst;
  a.order = m->order; a.done = m->done; a.ctl = m->ctl; a.ntiles = m->ntiles; a.tx = m->tx; a.ty = m->ty; a.tz = m->tz; a.st = st;
  a.epoch = ++m->epoch;
  const Geo g = geo(s);
  ProfScope ps(s, mode == 0 ? "k_mic_setup" : mode == 1 ? "k_mic_forward" : "k_mic_backward");
  const size_t sm = micSmemBytes<8>();
  if (mode == 0) k_mic_wave<0, 8><<<m->ctas, 128, sm, s->stream>>>(g, a);
  else if (mode == 1) k_mic_wave<1, 8><<<m->ctas, 128, sm, s->stream>>>(g, a);
  else k_mic_wave<2, 8><<<m->ctas, 128, sm, s->stream>>>(g, a);
  s->launches++;
}

} // namespace mf
