// noalloc.cu -- a reduction entry point that allocates nothing (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops, harnesses that forbid it).
//
// mf_dot_product_noalloc computes GridDotProduct (source/conjugategrad.cpp:175: sum over idx of the FLOAT product a[idx]*b[idx],
// accumulated in double) on caller-owned device arrays.  The scratch the two-stage reduction needs (one partial per block and a
// ticket counter) lives in a file-scope `static __device__` array of MF_NOALLOC_SCRATCH_BYTES (33 024 bytes), which every GPU
// that loads the library holds until the process exits.  The entry point calls no cudaMalloc / cudaFree / allocating routine; it
// enqueues one 4-byte memset and one kernel on the caller's stream and returns without synchronising.
//
// * Limit: n <= MF_NOALLOC_MAX_ELEMS (33 554 432 floats per input = 2 x 256^3); larger sizes are rejected with an error.
// * All calls on a device share the one scratch array, which keeps the last call's contents; the ticket counter is cleared by
//   every call.  Calls that OVERLAP on the GPU (different streams, or graph branches without an edge between them) race on the
//   scratch and give wrong sums: serialise them.
// * The summation order is fixed (block b sums elements [b*8192, (b+1)*8192) lane-strided, the last block adds the partials in
//   block order), so a given (a, b, n) always returns the same bits.
#include "common.cuh"
#include <mutex>

namespace mf {

constexpr int NOALLOC_THREADS = 256;
constexpr int NOALLOC_PER_BLOCK = 8192;                       // elements one block sums
constexpr int NOALLOC_MAX_BLOCKS = 4096;
constexpr long long NOALLOC_MAX_ELEMS = (long long)NOALLOC_MAX_BLOCKS * NOALLOC_PER_BLOCK;
constexpr size_t NOALLOC_SCRATCH_BYTES = sizeof(double) * NOALLOC_MAX_BLOCKS + 256;     // partials, then the counter's own 256-byte line
static_assert(MF_NOALLOC_MAX_ELEMS == NOALLOC_MAX_ELEMS, "include/mantacuda.h and noalloc.cu disagree on the limit");
static_assert(MF_NOALLOC_SCRATCH_BYTES == NOALLOC_SCRATCH_BYTES, "include/mantacuda.h and noalloc.cu disagree on the scratch size");

// In host code this name is only a stand-in for the symbol, not a device address: the host side asks the runtime for the
// address (noallocScratch below) and passes that to the kernel.
static __device__ __align__(256) unsigned char g_noallocScratch[NOALLOC_SCRATCH_BYTES];

__global__ void __launch_bounds__(NOALLOC_THREADS) k_dot_noalloc(const float* __restrict__ a, const float* __restrict__ b, long long n,
                                                                 double* __restrict__ out, unsigned char* scratch) {
  __shared__ double sh[NOALLOC_THREADS / 32];
  __shared__ bool isLast;
  double* partials = reinterpret_cast<double*>(scratch);
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch + sizeof(double) * NOALLOC_MAX_BLOCKS);
  const long long base = (long long)blockIdx.x * NOALLOC_PER_BLOCK;
  double acc = 0.0;
#pragma unroll 4
  for (int q = threadIdx.x; q < NOALLOC_PER_BLOCK; q += NOALLOC_THREADS) {
    const long long i = base + q;
    if (i < n) acc += (double)(a[i] * b[i]);      // float product, double sum (GridDotProduct)
  }
  acc = warpSumD(acc);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double v = 0.0;
    for (int w = 0; w < NOALLOC_THREADS / 32; w++) v += sh[w];
    partials[blockIdx.x] = v;
    __threadfence();
    isLast = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!isLast) return;
  __threadfence();
  // the last block: partials in block order (thread t takes b = t, t + 256, ...; then warp / block tree as above)
  double r = 0.0;
  for (unsigned int q = threadIdx.x; q < gridDim.x; q += NOALLOC_THREADS) r += ((volatile double*)partials)[q];
  r = warpSumD(r);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = r;
  __syncthreads();
  if (threadIdx.x == 0) {
    double v = 0.0;
    for (int w = 0; w < NOALLOC_THREADS / 32; w++) v += sh[w];
    *out = v;
    *counter = 0;
  }
}

// device address of the scratch array on the CURRENT device; asked from the runtime once per device
static unsigned char* noallocScratch() {
  static std::mutex mu;
  static unsigned char* ptrs[64] = {};
  int dev = -1;
  CUDA_CHECK(cudaGetDevice(&dev));
  mfAssert(dev >= 0 && dev < 64, "mf_dot_product_noalloc: unexpected device ordinal " << dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    unsigned char* p = nullptr;
    CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_noallocScratch));
    mfAssert(p != nullptr, "mf_dot_product_noalloc: cudaGetSymbolAddress returned a null scratch address");
    ptrs[dev] = p;
  }
  return ptrs[dev];
}

} // namespace mf

using namespace mf;

extern "C" int mf_dot_product_noalloc(const float* a_dev, const float* b_dev, long long n, double* out_dev, void* stream) {
  MF_API_BEGIN
  mfAssert(a_dev && b_dev && out_dev, "mf_dot_product_noalloc: null argument");
  mfAssert(n >= 0, "mf_dot_product_noalloc: negative size " << n);
  mfAssert(n <= NOALLOC_MAX_ELEMS, "mf_dot_product_noalloc: n = " << n << " exceeds the limit of " << NOALLOC_MAX_ELEMS
           << " elements (MF_NOALLOC_MAX_ELEMS; the scratch array is static, " << NOALLOC_SCRATCH_BYTES << " bytes)");
  unsigned char* scratch = noallocScratch();
  cudaStream_t st = (cudaStream_t)stream;
  // the counter must start at zero whatever an earlier (possibly aborted) call left behind
  CUDA_CHECK(cudaMemsetAsync(scratch + sizeof(double) * NOALLOC_MAX_BLOCKS, 0, sizeof(unsigned int), st));
  const unsigned int blocks = n == 0 ? 1u : (unsigned int)((n + NOALLOC_PER_BLOCK - 1) / NOALLOC_PER_BLOCK);
  k_dot_noalloc<<<blocks, NOALLOC_THREADS, 0, st>>>(a_dev, b_dev, n, out_dev, scratch);
  CUDA_CHECK(cudaGetLastError());
  MF_API_END
}
