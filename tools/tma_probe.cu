// tma_probe.cu -- development probe: one cp.async.bulk.tensor.3d box load per variant, checks the bytes that landed.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o build_tmp/tma_probe tools/tma_probe.cu && build_tmp/tma_probe
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)
static int hBX = 36, hBY = 20;
__constant__ int BX, BY;
__device__ __forceinline__ unsigned sa(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
struct Maps3 { CUtensorMap m[3]; };
template <int VARIANT>
__global__ void k(const __grid_constant__ CUtensorMap one, const __grid_constant__ Maps3 three, float* out, int x0, int y0, int z0) {
  extern __shared__ __align__(128) unsigned char sm[];
  float* box = reinterpret_cast<float*>(sm);
  unsigned long long* bar = reinterpret_cast<unsigned long long*>(sm + 4096);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sa(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sa(bar)), "r"(BX * BY * 4) : "memory");
    const CUtensorMap* mp = VARIANT == 0 ? &one : &three.m[1];
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(sa(box)), "l"(mp), "r"(x0), "r"(y0), "r"(z0), "r"(sa(bar)) : "memory");
  }
  unsigned ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(sa(bar)), "r"(0) : "memory");
  } while (!ok);
  for (int i = threadIdx.x; i < BX * BY; i += blockDim.x) out[i] = box[i];
}
int main(int argc, char** argv) {
  if (argc > 2) { hBX = atoi(argv[1]); hBY = atoi(argv[2]); }
  const int ax0 = argc > 3 ? atoi(argv[3]) : -2, ay0 = argc > 4 ? atoi(argv[4]) : 14;
  cudaMemcpyToSymbol(BX, &hBX, 4); cudaMemcpyToSymbol(BY, &hBY, 4);
  const int nx = 64, ny = 48, nz = 40;
  std::vector<float> h((size_t)nx * ny * nz);
  for (size_t i = 0; i < h.size(); i++) h[i] = (float)i;
  float *d, *out; CK(cudaMalloc(&d, h.size() * 4 * 3)); CK(cudaMalloc(&out, hBX * hBY * 4));
  for (int c = 0; c < 3; c++) CK(cudaMemcpy(d + c * h.size(), h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                         CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  Fn enc = (Fn)p;
  Maps3 three; CUtensorMap one;
  for (int c = 0; c < 3; c++) {
    const cuuint64_t dims[3] = { nx, ny, nz }; const cuuint64_t str[2] = { nx * 4, (cuuint64_t)nx * ny * 4 };
    const cuuint32_t box[3] = { (cuuint32_t)hBX, (cuuint32_t)hBY, 1 }; const cuuint32_t es[3] = { 1, 1, 1 };
    CUresult r = enc(&three.m[c], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d + c * h.size(), dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d -> %d\n", c, (int)r);
  }
  one = three.m[0];
  std::vector<float> o(hBX * hBY);
  for (int variant = 0; variant < 2; variant++) {
    const int x0 = ax0, y0 = ay0, z0 = 3;
    if (variant == 0) k<0><<<1, 128, 8192>>>(one, three, out, x0, y0, z0); else k<1><<<1, 128, 8192>>>(one, three, out, x0, y0, z0);
    cudaError_t e = cudaDeviceSynchronize();
    printf("variant %d: %s\n", variant, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    CK(cudaMemcpy(o.data(), out, o.size() * 4, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int r = 0; r < hBY; r++) for (int c = 0; c < hBX; c++) {
      const int x = x0 + c, y = y0 + r; const float want = (x < 0 || x >= nx || y < 0 || y >= ny) ? 0.f : (float)(x + nx * (y + ny * z0));
      if (o[r * hBX + c] != want) bad++;
    }
    printf("variant %d: %d wrong of %d\n", variant, bad, hBX * hBY);
  }
  return 0;
}
