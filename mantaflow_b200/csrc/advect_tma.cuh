// advect_tma.cuh -- MAC-grid advection with TMA-staged 3-D tiles (north star: "TMA-staged 3D tiles with halos in shared memory for
// advection"), the variant VERDICT r1 asked to be built and measured next to the per-thread-LDG kernels of advect_fast.cuh.
//
// A CTA owns a 32 x 16 tile in x / y and marches through a chunk of z.  Every plane of the three velocity components enters
// shared memory ONCE per CTA as a (32+8) x (16+4) box fetched by one `cp.async.bulk.tensor.3d` per component (one CUtensorMap per
// SoA plane, completion on an mbarrier); a ring of 8 plane slots per component keeps planes k-2 .. k+2 resident while the loads
// of the next three planes are in flight.  From shared memory, with immediate offsets:
//   * the face-velocity stencils getAtMACX/Y/Z (source/grid.h:473-506: 27 of the 51 loads per cell of k_sl_mac_f),
//   * the back-trace gather (source/util/interpol.h:72-94) whenever its 2 x 2 x 2 cell lies inside the staged window
//     (|v| dt < 2 cells: every trace of a CFL-limited step), with the per-thread global gather as the fallback otherwise.
// Same float operations in the same order as k_sl_mac_f / k_mc_mac_f: bit-identical results (tests/test_kernels_gpu.py).
// Out-of-range box elements are zero-filled by the TMA unit and never enter a result (border cells are written 0 like
// the reference's cleared temp grid, interior traces clamp their cell index into the grid).
#pragma once
#include <cuda.h>
#include "advect_fast.cuh"

namespace mf {

// the x halo is 4 cells, not 2: the first element of a box row must sit on a 16-byte boundary of global memory (a box that starts
// at x0 - 2 raises 'illegal instruction' on B200, tools/tma_probe.cu)
constexpr int TMA_TX = 32, TMA_TY = 16, TMA_H = 2, TMA_HX = 4, TMA_BX = TMA_TX + 2 * TMA_HX, TMA_BY = TMA_TY + 2 * TMA_H, TMA_R = 8;
constexpr int TMA_SLOT_BYTES = ((TMA_BX * TMA_BY * 4 + 127) / 128) * 128;     // 3200: TMA destinations are 128-byte aligned
constexpr int TMA_SLOT = TMA_SLOT_BYTES / 4;
constexpr int TMA_BOX_BYTES = TMA_BX * TMA_BY * 4;

__device__ __forceinline__ unsigned smemAddr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarInit(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbarExpectTx(unsigned long long* bar, int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(unsigned long long* bar, int parity) {
  unsigned ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(smemAddr(bar)), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void tmaLoad3D(void* dst, const CUtensorMap* map, int x, int y, int z, unsigned long long* bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
               ::"r"(smemAddr(dst)), "l"(map), "r"(x), "r"(y), "r"(z), "r"(smemAddr(bar)) : "memory");
}

struct TmaMaps3 { CUtensorMap m[3]; };

// one staged set of three component planes: ring[c][slot]
struct TmaRing {
  float* base; int zbase;
  __device__ __forceinline__ const float* plane(int c, int z) const { return base + (c * TMA_R + ((z - zbase) & (TMA_R - 1))) * TMA_SLOT; }
};

// getAtMACX/Y/Z times dt from the staged planes; o = offset of the cell inside a box (row * TMA_BX + column)
template <int C>
__device__ __forceinline__ void faceVelDtS(const TmaRing& r, int k, int o, float dt, float& vx, float& vy, float& vz) {
  const float* x = r.plane(0, k) + o; const float* y = r.plane(1, k) + o; const float* z = r.plane(2, k) + o;
  constexpr int Y = TMA_BX;
  float cx, cy, cz;
  if (C == 0) {
    const float* z1 = r.plane(2, k + 1) + o;
    cx = x[0];
    cy = 0.25f * (y[0] + y[-1] + y[Y] + y[Y - 1]);
    cz = 0.25f * (z[0] + z[-1] + z1[0] + z1[-1]);
  } else if (C == 1) {
    const float* z1 = r.plane(2, k + 1) + o;
    cx = 0.25f * (x[0] + x[-Y] + x[1] + x[1 - Y]);
    cy = y[0];
    cz = 0.25f * (z[0] + z[-Y] + z1[0] + z1[-Y]);
  } else {
    const float* xm = r.plane(0, k - 1) + o; const float* ym = r.plane(1, k - 1) + o;
    cx = 0.25f * (x[0] + xm[0] + x[1] + xm[1]);
    cy = 0.25f * (y[0] + ym[0] + y[Y] + ym[Y]);
    cz = z[0];
  }
  vx = cx * dt; vy = cy * dt; vz = cz * dt;
}

// interpol() of component plane c at (px, py, pz): from the staged window when the 2x2x2 cell lies inside it, else from global memory
__device__ __forceinline__ float interpolS(const TmaRing& r, int c, const float* __restrict__ glob, const GeoF& g, int x0, int y0, int k,
                                           float px, float py, float pz) {
  int xi, yi, zi; float s0, s1, t0, t1, f0, f1;
  axisW(px - 0.5f, g.nx, xi, s0, s1);
  axisW(py - 0.5f, g.ny, yi, t0, t1);
  axisW(pz - 0.5f, g.nz, zi, f0, f1);
  const int lx = xi - x0, ly = yi - y0;           // x0 / y0: grid coordinates of the box origin
  const float *p, *q; int Y;
  if ((unsigned)lx <= (unsigned)(TMA_BX - 2) && (unsigned)ly <= (unsigned)(TMA_BY - 2) && (unsigned)(zi - (k - TMA_H)) <= (unsigned)(2 * TMA_H - 1)) {
    const int o = ly * TMA_BX + lx;
    p = r.plane(c, zi) + o; q = r.plane(c, zi + 1) + o; Y = TMA_BX;
  } else {
    p = glob + (xi + g.nx * yi + g.Z * zi); q = p + g.Z; Y = g.nx;
  }
  return ((p[0] * t0 + p[Y] * t1) * s0 + (p[1] * t0 + p[Y + 1] * t1) * s1) * f0
       + ((q[0] * t0 + q[Y] * t1) * s0 + (q[1] * t0 + q[Y + 1] * t1) * s1) * f1;
}

// SemiLagrangeMAC (advection.cpp:45) for self-advection (src == vel): dst = interpolComponent(vel, pos - getAtMAC(vel) * dt)
__global__ void __launch_bounds__(256) k_sl_mac_tma(Geo gg, const __grid_constant__ TmaMaps3 maps, CV3 vel, V3 dst, float dt, int zc) {
  extern __shared__ __align__(128) unsigned char tmaSmem[];
  float* ring = reinterpret_cast<float*>(tmaSmem);
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(tmaSmem + 3 * TMA_R * TMA_SLOT_BYTES);
  const GeoF g = geoF(gg);
  const int tilesX = (g.nx + TMA_TX - 1) / TMA_TX, tilesY = (g.ny + TMA_TY - 1) / TMA_TY;
  int b = blockIdx.x;
  const int bx = b % tilesX; b /= tilesX;
  const int by = b % tilesY; b /= tilesY;
  const int kz0 = g.k0 + b * zc, kz1 = min(kz0 + zc, g.k1);
  const int x0 = bx * TMA_TX - TMA_HX, y0 = by * TMA_TY - TMA_H;
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  TmaRing r; r.base = ring; r.zbase = kz0 - TMA_H;
  if (tid == 0) {
    for (int s = 0; s < TMA_R; s++) mbarInit(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int zfirst = kz0 - TMA_H, zlast = kz1 - 1 + TMA_H;      // planes this CTA stages
  auto issue = [&](int z) {                                     // thread 0: the three component planes of z
    const int s = (z - r.zbase) & (TMA_R - 1);
    mbarExpectTx(&bars[s], 3 * TMA_BOX_BYTES);
#pragma unroll
    for (int c = 0; c < 3; c++) tmaLoad3D(ring + (c * TMA_R + s) * TMA_SLOT, &maps.m[c], x0, y0, z, &bars[s]);
  };
  if (tid == 0) for (int z = zfirst; z <= min(zlast, zfirst + TMA_R - 1); z++) issue(z);
  const int i = bx * TMA_TX + tx;
  for (int k = kz0; k < kz1; k++) {
    // planes k-2 .. k+2 are needed; k+2 is the newest: wait for every slot not yet seen (the first plane waits for five)
    for (int z = (k == kz0 ? k - TMA_H : k + TMA_H); z <= k + TMA_H; z++) {
      const int n = z - r.zbase;
      mbarWait(&bars[n & (TMA_R - 1)], (n / TMA_R) & 1);
    }
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int j = by * TMA_TY + ty + 8 * h;
      if (i < g.nx && j < g.ny) {
        const int idx = i + g.nx * (j + g.ny * k);
        const bool border = i < 1 || j < 1 || k < 1 || i >= g.nx - 1 || j >= g.ny - 1 || k >= g.nz - 1;
        float rx = 0.0f, ry = 0.0f, rz = 0.0f;
        if (!border) {
          const int o = (ty + 8 * h + TMA_H) * TMA_BX + tx + TMA_HX;
          float vx, vy, vz;
          faceVelDtS<0>(r, k, o, dt, vx, vy, vz);
          rx = interpolS(r, 0, vel.x, g, x0, y0, k, (i + 0.5f) - vx, (j + 0.5f) - vy, (k + 0.5f) - vz);
          faceVelDtS<1>(r, k, o, dt, vx, vy, vz);
          ry = interpolS(r, 1, vel.y, g, x0, y0, k, (i + 0.5f) - vx, (j + 0.5f) - vy, (k + 0.5f) - vz);
          faceVelDtS<2>(r, k, o, dt, vx, vy, vz);
          rz = interpolS(r, 2, vel.z, g, x0, y0, k, (i + 0.5f) - vx, (j + 0.5f) - vy, (k + 0.5f) - vz);
        }
        dst.x[idx] = rx; dst.y[idx] = ry; dst.z[idx] = rz;
      }
    }
    __syncthreads();                               // every thread is done with plane k-2: its slot takes plane k-2+R
    if (tid == 0 && k - TMA_H + TMA_R <= zlast) issue(k - TMA_H + TMA_R);
  }
}

inline size_t tmaSmemBytes() { return (size_t)3 * TMA_R * TMA_SLOT_BYTES + TMA_R * sizeof(unsigned long long); }

// CUtensorMap of one SoA plane: 3-D float tensor (nx, ny, nz), box (TMA_BX, TMA_BY, 1), no swizzle, zero fill outside
typedef CUresult (*TensorMapEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                           const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline CUtensorMap tmaPlaneMap(const Solver* s, const float* plane) {
  static TensorMapEncodeTiledFn encode = nullptr;
  if (!encode) {
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    mfAssert(p && q == cudaDriverEntryPointSuccess, "mantacuda: cuTensorMapEncodeTiled not available");
    encode = reinterpret_cast<TensorMapEncodeTiledFn>(p);
  }
  CUtensorMap m;
  const cuuint64_t dims[3] = { (cuuint64_t)s->nx, (cuuint64_t)s->ny, (cuuint64_t)s->nz };
  const cuuint64_t strides[2] = { (cuuint64_t)s->nx * 4, (cuuint64_t)s->nx * s->ny * 4 };
  const cuuint32_t box[3] = { TMA_BX, TMA_BY, 1 };
  const cuuint32_t estr[3] = { 1, 1, 1 };
  const CUresult r = encode(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(plane), dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  mfAssert(r == CUDA_SUCCESS, "mantacuda: cuTensorMapEncodeTiled failed (" << (int)r << ")");
  return m;
}

// usable: single GPU, 3-D, rows a multiple of 16 bytes
inline bool tmaUsable(const Solver* s) {
  return s->is3D && !s->multi() && (s->nx % 4 == 0) && s->nx >= TMA_TX && s->ny >= TMA_TY && s->nz >= 4;
}

} // namespace mf
