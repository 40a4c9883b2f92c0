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

template <int TZ> inline size_t micSmemBytes() { return sizeof(float) * 5 * (size_t)(19 * 18 * (TZ + 1)); }

inline void micDestroy(MicPlan* m) {
  if (!m) return;
  cudaFree(m->order); cudaFree(m->done); cudaFree(m->ctl);
  delete m;
}

inline MicPlan* micPlan(Solver* s) {
  if (s->mic) return s->mic;
  MicPlan* m = new MicPlan();
  m->tzEdge = getenv("MF_MIC_TZ") ? atoi(getenv("MF_MIC_TZ")) : 8;
  mfAssert(m->tzEdge == 8 || m->tzEdge == 16, "MF_MIC_TZ must be 8 or 16");
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
  int occ = 1;
  if (m->tzEdge == 8) {
    CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<0, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<8>()));
    CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<1, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<8>()));
    CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<2, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<8>()));
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_mic_wave<1, 8>, 128, micSmemBytes<8>()));
  } else {
    CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<0, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<16>()));
    CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<1, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<16>()));
    CUDA_CHECK(cudaFuncSetAttribute(k_mic_wave<2, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)micSmemBytes<16>()));
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_mic_wave<1, 16>, 256, micSmemBytes<16>()));
  }
  if (occ < 1) occ = 1;
  m->ctas = s->numSMs * occ; if (m->ctas > m->ntiles) m->ctas = m->ntiles;
  s->mic = m;
  return m;
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
  if (m->tzEdge == 8) {
    const size_t sm = micSmemBytes<8>();
    if (mode == 0) k_mic_wave<0, 8><<<m->ctas, 128, sm, s->stream>>>(g, a);
    else if (mode == 1) k_mic_wave<1, 8><<<m->ctas, 128, sm, s->stream>>>(g, a);
    else k_mic_wave<2, 8><<<m->ctas, 128, sm, s->stream>>>(g, a);
  } else {
    const size_t sm = micSmemBytes<16>();
    if (mode == 0) k_mic_wave<0, 16><<<m->ctas, 256, sm, s->stream>>>(g, a);
    else if (mode == 1) k_mic_wave<1, 16><<<m->ctas, 256, sm, s->stream>>>(g, a);
    else k_mic_wave<2, 16><<<m->ctas, 256, sm, s->stream>>>(g, a);
  }
  s->launches++;
}

} // namespace mf
