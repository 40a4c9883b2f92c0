// barrier_probe.cu -- cost of a software grid barrier on one GPU (development probe for mg_tail.cuh).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/barrier_probe tools/barrier_probe.cu && /tmp/barrier_probe
// Variants: 0 red.release + ld.acquire spin by thread 0 (what k_mg_tail uses); 1 the same with __threadfence on both sides;
// 2 relaxed atomicAdd + volatile spin + fences (cooperative-groups style); 3 variant 0 with a per-CTA flag array written by the
// last arriver (each CTA polls its own line).
#include <cstdio>
#include <cuda_runtime.h>

template <int V>
__global__ void __launch_bounds__(512, 1) k_probe(unsigned* bar, unsigned* flags, int iters, float* sink, int work) {
  unsigned target = 0;
  float acc = 0.f;
  for (int it = 0; it < iters; it++) {
    for (int w = 0; w < work; w++) acc += __ldcg(sink + ((threadIdx.x + blockIdx.x * blockDim.x + w * 4099) & 0xffff));
    __syncthreads();
    if (threadIdx.x == 0) {
      target += gridDim.x;
      if (V == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(bar) : "memory");
        unsigned v;
        do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory"); } while ((int)(v - target) < 0);
      } else if (V == 1) {
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(bar) : "memory");
        unsigned v;
        do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory"); } while ((int)(v - target) < 0);
        __threadfence();
      } else if (V == 2) {
        __threadfence();
        atomicAdd(bar, 1u);
        while ((int)(*(volatile unsigned*)bar - target) < 0) {}
        __threadfence();
      } else {
        unsigned old;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(bar) : "memory");
        if (old + 1 == target) {
          for (unsigned c = 0; c < gridDim.x; c++) asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(flags + 32 * c), "r"(target) : "memory");
        }
        unsigned v;
        do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + 32 * blockIdx.x) : "memory"); } while ((int)(v - target) < 0);
      }
    }
    __syncthreads();
  }
  if (acc == 123.456f) sink[0] = acc;
}

template <int V>
static void run(int G, int T, int iters, int work, unsigned* bar, unsigned* flags, float* sink) {
  cudaMemset(bar, 0, 4); cudaMemset(flags, 0, 4 * 32 * 1024);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k_probe<V><<<G, T>>>(bar, flags, 10, sink, work);
  cudaMemset(bar, 0, 4); cudaMemset(flags, 0, 4 * 32 * 1024);
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  k_probe<V><<<G, T>>>(bar, flags, iters, sink, work);
  cudaEventRecord(b);
  cudaError_t e = cudaDeviceSynchronize();
  float ms = 0; cudaEventElapsedTime(&ms, a, b);
  printf("variant %d  G=%3d T=%3d work=%2d : %.3f us per barrier+work  (%s)\n", V, G, T, work, 1e3 * ms / iters, cudaGetErrorString(e));
}

int main() {
  unsigned *bar, *flags; float* sink;
  cudaMalloc(&bar, 4); cudaMalloc(&flags, 4 * 32 * 1024); cudaMalloc(&sink, 4 * 65536); cudaMemset(sink, 0, 4 * 65536);
  const int iters = 2000;
  for (int G : {16, 74, 148}) {
    for (int work : {0, 1}) {
      run<0>(G, 512, iters, work, bar, flags, sink);
      run<1>(G, 512, iters, work, bar, flags, sink);
      run<2>(G, 512, iters, work, bar, flags, sink);
      run<3>(G, 512, iters, work, bar, flags, sink);
    }
  }
  run<0>(148, 128, iters, 0, bar, flags, sink);
  run<0>(296, 256, iters, 0, bar, flags, sink);
  return 0;
}
