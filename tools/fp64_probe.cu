// fp64_probe.cu -- what one iteration of the coarsest-level CG is made of, in SM cycles (development probe for k_mg_coarse_cg,
// DESIGN.md section 5 (iii)).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o /tmp/fp64_probe tools/fp64_probe.cu && /tmp/fp64_probe
// One CTA; every section is timed with clock64 by thread 0 between block barriers, REP repetitions.
// Measured on B200 at 736 threads (cycles per repetition): dependent DADD / DMUL / DFMA 12, double division 171, square root 177,
// warp tree sum of a double 237, block barrier 58, block sum with two barriers and one division by warp 0: 853, block sum with one
// barrier that every thread finishes itself: 1291 (the redundant divisions saturate the FP64 pipe).  At 32 threads: 8 / 129 /
// 106 / 185 / 14 / 738 / 372.
#include <cstdio>
#include <cuda_runtime.h>

constexpr int REP = 64;

__device__ __forceinline__ double warpSum(double a) {
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  return a;
}

__global__ void __launch_bounds__(1024) k_probe(double* io, long long* cyc) {
  __shared__ double part[32];
  __shared__ double res;
  const int t = threadIdx.x;
  double x = io[t], y = io[t + 1024] + 1.5;
  long long t0, t1;
  int sec = 0;
#define BEGIN __syncthreads(); t0 = clock64();
#define END   __syncthreads(); t1 = clock64(); if (t == 0) cyc[sec] = t1 - t0; sec++;

  BEGIN END                                                                   // 0: empty section (two barriers)
  BEGIN for (int i = 0; i < REP; i++) x = x + y; END                          // 1: dependent DADD
  BEGIN for (int i = 0; i < REP; i++) x = x * y; END                          // 2: dependent DMUL
  BEGIN for (int i = 0; i < REP; i++) x = __fma_rn(x, y, y); END              // 3: dependent DFMA
  BEGIN for (int i = 0; i < REP; i++) x = y / (x + 2.0); END                  // 4: dependent divisions
  BEGIN for (int i = 0; i < REP; i++) x = sqrt(x * x + 2.0); END              // 5: dependent square roots
  BEGIN for (int i = 0; i < REP; i++) x = warpSum(x) * 0.03125; END           // 6: warp tree sums
  BEGIN for (int i = 0; i < REP; i++) __syncthreads(); END                    // 7: block barriers
  // 8: block sums as the kernel does them (tree, partials, barrier, warp 0 tree + division, barrier)
  BEGIN
  for (int i = 0; i < REP; i++) {
    double a = warpSum(x);
    if ((t & 31) == 0) part[t >> 5] = a;
    __syncthreads();
    if (t < 32) {
      a = (t < (int)(blockDim.x >> 5)) ? part[t] : 0.0;
      a = warpSum(a);
      if (t == 0) res = y / a;
    }
    __syncthreads();
    x = x * res + 1.0;
  }
  END
  // 9: block sums with one barrier (+ one that protects the single buffer), every thread adds the partials itself
  BEGIN
  for (int i = 0; i < REP; i++) {
    double a = warpSum(x);
    if ((t & 31) == 0) part[t >> 5] = a;
    __syncthreads();
    a = part[0];
    for (int k = 1; k < (int)(blockDim.x >> 5); k++) a += part[k];
    x = x * (y / a) + 1.0;
    __syncthreads();
  }
  END
  // 10: the vector update of an iteration with one division per thread
  BEGIN for (int i = 0; i < REP; i++) { x = x + y * 0.5; y = y - x * 0.25; x = y / (x + 3.0); } END
  io[t] = x + y;
}

int main() {
  double* io; long long* cyc;
  cudaMalloc(&io, 8 * 4096); cudaMalloc(&cyc, 8 * 32);
  double h[4096]; for (int i = 0; i < 4096; i++) h[i] = 1.0 + 1e-3 * i;
  const char* name[] = {"empty (2 barriers)", "dependent DADD", "dependent DMUL", "dependent DFMA", "dependent DDIV(+DADD)", "dependent DSQRT(+DFMA)",
                        "warp tree sum of a double", "block barrier", "block sum, 2 barriers, warp 0 finishes", "block sum, 1(+1) barrier, all threads finish",
                        "vector update with one division"};
  for (int nt : {32, 256, 736, 1024}) {
    cudaMemcpy(io, h, sizeof(h), cudaMemcpyHostToDevice);
    k_probe<<<1, nt>>>(io, cyc);
    k_probe<<<1, nt>>>(io, cyc);
    if (cudaGetLastError() != cudaSuccess) { printf("launch failed at %d threads\n", nt); continue; }
    long long c[32];
    cudaError_t e = cudaMemcpy(c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
    printf("--- %d threads (%s)\n", nt, cudaGetErrorString(e));
    for (int s = 0; s < 11; s++) printf("  %-46s %8.1f cycles per repetition\n", name[s], (double)(c[s] - (s ? c[0] : 0)) / (s ? REP : 1));
  }
  return 0;
}
