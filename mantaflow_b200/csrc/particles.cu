// particles.cu -- FLIP particle system on the device, SURVEY 8(f) rank 3.
// Reference: BasicParticleSystem / ParticleSystem<S> (source/particle.h:32-180, :198), ParticleDataImpl (source/particle.h:340+,
// particle.cpp:349-369), GridAdvectKernel / KnClampPositions / KnDeleteInObstacle / advectInGrid (particle.h:457-550),
// integratePointSet (source/util/integrator.h:27-78), sampleFlagsWithParticles (source/plugin/flip.cpp:32-56), markFluidCells
// (:143-188), knMapLinearVec3ToMACGrid / mapPartsToMAC (:619-661) with setInterpolMAC (source/util/interpol.h:166-213),
// mapMACToParts (:709-721), flipVelocityUpdate (:724-742).
//
// Positions, flags and particle data are SoA arrays in device memory.  Particle -> grid transfer: the reference scatters serially
// in particle order (KERNEL(pts, single)), so every face value is a float sum in particle order.  Here each particle emits its 8
// (weight, weight * value) contributions per component in the reference's order, the contributions are stably sorted by target
// face (cub radix sort) and one thread per face adds its run front to back -- the same sums, bit for bit, at any thread count.
// On z-slab partitioned solvers the particle arrays are replicated on every rank; the rank that owns slab 0 writes the
// particle -> grid transfers for the whole grid and every rank then passes rankSync().
#include "common.cuh"
#include "sample.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cmath>
#include <random>

namespace mf {

enum { PNEW = 1, PDELETE = 1 << 10 };            // ParticleBase::PNEW / PDELETE (particle.h:36, :41)

struct PData { int ncomp; float* d; };           // ncomp 1 (Real), 3 (Vec3, SoA planes of cap entries) or -1 (int)
struct Parts {
  Solver* s;
  long long n, cap;
  float* pos;                                    // x[cap] y[cap] z[cap]
  int* flag;
  std::vector<PData> pdata;
  // scratch of the integrators / transfers (grown on demand)
  float* scratch; size_t scratchBytes;
  float* px() const { return pos; } float* py() const { return pos + cap; } float* pz() const { return pos + 2 * cap; }
};

static void* scratchOf(Parts* p, size_t bytes) {
  if (bytes > p->scratchBytes) {
    if (p->scratch) CUDA_CHECK(cudaFree(p->scratch));
    CUDA_CHECK(cudaMalloc(&p->scratch, bytes)); p->scratchBytes = bytes;
  }
  return p->scratch;
}
static void reserve(Parts* p, long long cap) {
  if (cap <= p->cap) return;
  Solver* s = p->s;
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  float* np = nullptr; int* nf = nullptr;
  CUDA_CHECK(cudaMalloc(&np, sizeof(float) * 3 * (size_t)cap)); CUDA_CHECK(cudaMalloc(&nf, sizeof(int) * (size_t)cap));
  for (int c = 0; c < 3; c++) if (p->n) CUDA_CHECK(cudaMemcpy(np + c * cap, p->pos + c * p->cap, sizeof(float) * (size_t)p->n, cudaMemcpyDeviceToDevice));
  if (p->n) CUDA_CHECK(cudaMemcpy(nf, p->flag, sizeof(int) * (size_t)p->n, cudaMemcpyDeviceToDevice));
  cudaFree(p->pos); cudaFree(p->flag); p->pos = np; p->flag = nf;
  for (PData& d : p->pdata) {
    const int nc = d.ncomp == 3 ? 3 : 1;
    float* nd = nullptr;
    CUDA_CHECK(cudaMalloc(&nd, sizeof(float) * nc * (size_t)cap));
    CUDA_CHECK(cudaMemset(nd, 0, sizeof(float) * nc * (size_t)cap));
    for (int c = 0; c < nc; c++) if (p->n) CUDA_CHECK(cudaMemcpy(nd + c * cap, d.d + c * p->cap, sizeof(float) * (size_t)p->n, cudaMemcpyDeviceToDevice));
    cudaFree(d.d); d.d = nd;
  }
  p->cap = cap;
}

// ---- device helpers --------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool inBoundsI(const Geo& g, int x, int y, int z, int bnd) {      // GridBase::isInBounds(Vec3i, bnd), grid.h:430
  bool r = x >= bnd && y >= bnd && x < g.nx - bnd && y < g.ny - bnd;
  if (g.is3D) r = r && (z >= bnd && z < g.nz - bnd); else r = r && (z == 0);
  return r;
}
__device__ __forceinline__ int flagAt(const Geo& g, const int* __restrict__ flags, float x, float y, float z) {   // FlagGrid::getAt(Vec3), grid.h:323
  return flags[(long long)(int)x + (long long)g.nx * (int)y + g.Z * (int)z];
}

// GridAdvectKernel (particle.h:457-481): u = vel(pos) * dt, or 0 for deleted / stopped particles
__global__ void k_part_velocity(Geo g, long long n, const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz,
                                int* __restrict__ pflag, CV3 vel, const int* __restrict__ flags, float dt, int deleteInObstacle, int stopInObstacle,
                                int skipNew, float* __restrict__ ux, float* __restrict__ uy, float* __restrict__ uz) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int f = pflag[t];
  if ((f & PDELETE) || (skipNew && (f & PNEW))) { ux[t] = 0.f; uy[t] = 0.f; uz[t] = 0.f; return; }
  const float x = px[t], y = py[t], z = pz[t];
  if (deleteInObstacle || stopInObstacle) {
    if (!inBoundsI(g, (int)x, (int)y, (int)z, 1) || fIsObstacle(flagAt(g, flags, x, y, z))) {
      if (stopInObstacle) { ux[t] = 0.f; uy[t] = 0.f; uz[t] = 0.f; }     // (otherwise u keeps its previous value, as in the reference)
      if (deleteInObstacle) pflag[t] = f | PDELETE;
      return;
    }
  }
  float vx, vy, vz;
  interpolMAC(vel.x, vel.y, vel.z, g, x, y, z, vx, vy, vz);
  ux[t] = vx * dt; uy[t] = vy * dt; uz[t] = vz * dt;
}
// integratePointSet stages (integrator.h:33-72): x = x0 + a * u ; total (+)= b * u.  Products with the double literal 0.5 and the
// int 2 are exact in float.
__global__ void k_part_stage(long long n, float* __restrict__ x, const float* __restrict__ x0, const float* __restrict__ u, float a,
                             float* __restrict__ total, float b, int initTotal) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const float ut = u[t];
  x[t] = x0[t] + a * ut;
  if (total) {
    // RK4 in this fork starts the accumulator at u and adds u once more before the first stage (integrator.h:51-56)
    float tt = initTotal ? ut : total[t];
    tt += b * ut;
    total[t] = tt;
  }
}
__global__ void k_part_rk4_final(long long n, float* __restrict__ x, const float* __restrict__ x0, const float* __restrict__ u,
                                 const float* __restrict__ total) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  x[t] = x0[t] + (float)(1. / 6.) * (total[t] + u[t]);
}

// bisectBacktracePos (particle.h:494-505)
__device__ __forceinline__ void bisectBack(const Geo& g, const int* __restrict__ flags, float ox, float oy, float oz, float& nx, float& ny, float& nz) {
  float s = 0.f;
  for (int i = 1; i < 5; ++i) {
    const float ds = (float)(1. / (double)(float)(1 << i));
    const float sn = s + ds;
    const double so = 1. - (double)sn;
    const float tx = (float)((double)ox * so) + nx * sn, ty = (float)((double)oy * so) + ny * sn, tz = (float)((double)oz * so) + nz * sn;
    if (!fIsObstacle(flagAt(g, flags, tx, ty, tz))) s += ds;
  }
  const double so = 1. - (double)s;
  nx = (float)((double)ox * so) + nx * s; ny = (float)((double)oy * so) + ny * s; nz = (float)((double)oz * so) + nz * s;
}
// KnClampPositions (particle.h:507-523)
__global__ void k_part_clamp(Geo g, long long n, float* __restrict__ px, float* __restrict__ py, float* __restrict__ pz, const int* __restrict__ pflag,
                             const int* __restrict__ flags, const float* __restrict__ ox, const float* __restrict__ oy, const float* __restrict__ oz,
                             int stopInObstacle) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  if (pflag[t] & PDELETE) return;
  float x = px[t], y = py[t], z = pz[t];
  if (!inBoundsI(g, (int)x, (int)y, (int)z, 0)) {
    const float hx = (float)g.nx - 1.f, hy = (float)g.ny - 1.f, hz = (float)g.nz - 1.f;
    x = (x < 0.f) ? 0.f : ((x > hx) ? hx : x);
    y = (y < 0.f) ? 0.f : ((y > hy) ? hy : y);
    z = (z < 0.f) ? 0.f : ((z > hz) ? hz : z);
  }
  if (stopInObstacle && fIsObstacle(flagAt(g, flags, x, y, z))) bisectBack(g, flags, ox[t], oy[t], oz[t], x, y, z);
  px[t] = x; py[t] = y; pz[t] = z;
}
// KnDeleteInObstacle (particle.h:485-491)
__global__ void k_part_delete_in_obstacle(Geo g, long long n, const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz,
                                          int* __restrict__ pflag, const int* __restrict__ flags) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int f = pflag[t];
  if (f & PDELETE) return;
  const float x = px[t], y = py[t], z = pz[t];
  if (!inBoundsI(g, (int)x, (int)y, (int)z, 1) || fIsObstacle(flagAt(g, flags, x, y, z))) pflag[t] = f | PDELETE;
}

// markFluidCells (flip.cpp:143-147, :166-180)
// whole-array fills / copies as kernels: on a partitioned solver the arrays span the memory of several GPUs
__global__ void k_zero_words(long long n, unsigned* __restrict__ d) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) d[i] = 0u;
}
__global__ void k_copy_words(long long n, unsigned* __restrict__ d, const unsigned* __restrict__ src) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) d[i] = src[i];
}
__global__ void k_clear_fluid_flags(long long n, int* __restrict__ flags) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int f = flags[t];
  if (f & MF_FLUID) flags[t] = (f | MF_EMPTY) & ~MF_FLUID;
}
__global__ void k_mark_fluid(Geo g, long long n, const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz,
                             const int* __restrict__ pflag, int* flags) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  if (pflag[t] & PDELETE) return;
  const int x = (int)px[t], y = (int)py[t], z = (int)pz[t];
  if (!(x >= 0 && y >= 0 && z >= 0 && x < g.nx && y < g.ny && z < g.nz)) return;
  const long long c = (long long)x + (long long)g.nx * y + g.Z * z;
  // several particles of a cell write the same value; the test reads either the old or the new flag, both give this value
  const int f = flags[c];
  if (f & MF_EMPTY) flags[c] = (f | MF_FLUID) & ~MF_EMPTY;
}
// knSetNbObstacle (flip.cpp:149-164), out of place
__global__ void __launch_bounds__(256) k_set_nb_obstacle(Geo g, int* __restrict__ nflags, const int* __restrict__ flags, const float* __restrict__ phiObs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  if (i < 1 || j < 1 || i >= g.nx - 1 || j >= g.ny - 1 || (g.is3D && (k < 1 || k >= g.nz - 1))) return;
  const long long idx = (long long)i + (long long)g.nx * j + g.Z * k, Y = g.nx, Z = g.Z;
  if (phiObs[idx] > 0.f) return;
  const int f = flags[idx];
  if (!(f & MF_EMPTY)) return;
  bool set = false;
  if (fIsFluid(flags[idx - 1]) && phiObs[idx + 1] <= 0.f) set = true;
  if (fIsFluid(flags[idx + 1]) && phiObs[idx - 1] <= 0.f) set = true;
  if (fIsFluid(flags[idx - Y]) && phiObs[idx + Y] <= 0.f) set = true;
  if (fIsFluid(flags[idx + Y]) && phiObs[idx - Y] <= 0.f) set = true;
  if (g.is3D) {
    if (fIsFluid(flags[idx - Z]) && phiObs[idx + Z] <= 0.f) set = true;
    if (fIsFluid(flags[idx + Z]) && phiObs[idx - Z] <= 0.f) set = true;
  }
  if (set) nflags[idx] = (f | MF_FLUID) & ~MF_EMPTY;
}

// knMapLinearMACGridToVec3_PIC / _FLIP (flip.cpp:709-736)
__global__ void k_grid_to_parts(Geo g, long long n, const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz,
                                const int* __restrict__ pflag, CV3 vel, CV3 velOld, int flip, float flipRatio,
                                float* __restrict__ vx, float* __restrict__ vy, float* __restrict__ vz) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  if (pflag[t] & PDELETE) return;
  const float x = px[t], y = py[t], z = pz[t];
  float ax, ay, az;
  interpolMAC(vel.x, vel.y, vel.z, g, x, y, z, ax, ay, az);                       // v2
  if (!flip) { vx[t] = ax; vy[t] = ay; vz[t] = az; return; }
  float bx, by, bz;
  interpolMAC(velOld.x, velOld.y, velOld.z, g, x, y, z, bx, by, bz);              // v1
  const double om = 1.0 - (double)flipRatio;
  vx[t] = flipRatio * (vx[t] + (ax - bx)) + (float)(om * (double)ax);
  vy[t] = flipRatio * (vy[t] + (ay - by)) + (float)(om * (double)ay);
  vz[t] = flipRatio * (vz[t] + (az - bz)) + (float)(om * (double)az);
}

// ---- particle -> MAC grid (setInterpolMAC, interpol.h:166-213) ------------------------------------------------------------
// contribution slot order inside one particle and component, as the reference adds them.  X, Y: the four z+1 corners first.
//   slot -> (dx, dy, dz):  X/Y: (0,0,1) (1,0,1) (0,1,1) (1,1,1) (0,0,0) (1,0,0) (0,1,0) (1,1,0);  Z: the z corners first
template <int C>
__global__ void k_part_contrib(Geo g, long long n, const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz,
                               const int* __restrict__ pflag, const float* __restrict__ val, unsigned* __restrict__ keys, unsigned* __restrict__ ids,
                               float* __restrict__ wOut, float* __restrict__ wvOut) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  const long long o = t * 8;
  if (pflag[t] & PDELETE) {
#pragma unroll
    for (int q = 0; q < 8; q++) { keys[o + q] = 0xffffffffu; ids[o + q] = (unsigned)(o + q); wOut[o + q] = 0.f; wvOut[o + q] = 0.f; }
    return;
  }
  const float posx = px[t], posy = py[t], posz = pz[t];
  const IdxW w = buildIndex(g, posx, posy, posz);
  // the component's own axis uses the unshifted index (BUILD_INDEX_SHIFT, interpol.h:116-129)
  const float p = (C == 0) ? posx : (C == 1 ? posy : posz);
  const int nn = (C == 0) ? g.nx : (C == 1 ? g.ny : g.nz);
  int si = (int)p; float s1 = p - (float)si, s0 = 1.0f - s1;
  if (p < 0.f) { si = 0; s0 = 1.0f; s1 = 0.0f; }
  if (C < 2 || g.nz > 1) { if (si >= nn - 1) { si = nn - 2; s0 = 0.0f; s1 = 1.0f; } }
  const int xi = (C == 0) ? si : w.xi, yi = (C == 1) ? si : w.yi, zi = (C == 2) ? si : w.zi;
  const float a0 = (C == 0) ? s0 : w.s0, a1 = (C == 0) ? s1 : w.s1;      // x weights
  const float b0 = (C == 1) ? s0 : w.t0, b1 = (C == 1) ? s1 : w.t1;      // y weights
  const float c0 = (C == 2) ? s0 : w.f0, c1 = (C == 2) ? s1 : w.f1;      // z weights
  const float s0f0 = a0 * c0, s1f0 = a1 * c0, s0f1 = a0 * c1, s1f1 = a1 * c1;
  const float ww[8] = { b0 * s0f1, b0 * s1f1, b1 * s0f1, b1 * s1f1, b0 * s0f0, b0 * s1f0, b1 * s0f0, b1 * s1f0 };   // wz wxz wyz wxyz w0 wx wy wxy
  const long long base = ((long long)zi * g.ny + yi) * g.nx + xi;
  const float v = val[t];
#pragma unroll
  for (int q = 0; q < 8; q++) {
    // X and Y add the z+1 corners first, Z the z corners first
    const int e = (C == 2) ? (q ^ 4) : q;
    const int dz = (e < 4) ? 1 : 0, dy = (e >> 1) & 1, dx = e & 1;
    const long long cell = base + dx + (long long)g.nx * dy + g.Z * dz;
    keys[o + q] = (unsigned)cell; ids[o + q] = (unsigned)(o + q);
    wOut[o + q] = ww[e]; wvOut[o + q] = ww[e] * v;
  }
}
// one thread per run of equal keys: front-to-back float sums into the (cleared) grids
__global__ void k_part_accumulate(long long m, const unsigned* __restrict__ keys, const unsigned* __restrict__ ids, const float* __restrict__ w,
                                  const float* __restrict__ wv, float* __restrict__ sumGrid, float* __restrict__ refGrid) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= m) return;
  const unsigned key = keys[t];
  if (key == 0xffffffffu || (t > 0 && keys[t - 1] == key)) return;
  float s = 0.f, r = 0.f;
  for (long long q = t; q < m && keys[q] == key; q++) { const unsigned id = ids[q]; s += w[id]; r += wv[id]; }
  sumGrid[key] = s; refGrid[key] = r;
}
// weight->stomp(VECTOR_EPSILON); vel.safeDivide(weight)  (grid.cpp:247-249, general.h:150)
__global__ void k_stomp_divide(long long n3, float* __restrict__ vel, float* __restrict__ weight) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n3) return;
  float wgt = weight[t];
  if (wgt < 1e-6f) { wgt = 0.f; weight[t] = 0.f; }
  const float a = vel[t];
  vel[t] = (wgt != 0.f) ? a / wgt : a;
}

static inline unsigned pblocks(long long n) { return cdiv(n > 0 ? n : 1, 256); }
static void check_api(int rc) { if (rc) throw std::runtime_error(mf_last_error()); }

static Parts* P(mf_parts* p) { return reinterpret_cast<Parts*>(p); }
static const Parts* P(const mf_parts* p) { return reinterpret_cast<const Parts*>(p); }
static const PData& pdataOf(const Parts* p, int pd, int ncomp, const char* fn) {
  mfAssert(pd >= 0 && pd < (int)p->pdata.size() && p->pdata[pd].ncomp == ncomp, fn << ": particle data of the wrong type or missing");
  return p->pdata[pd];
}

} // namespace mf

using namespace mf;

extern "C" {

mf_parts* mf_parts_create(mf_solver* sp) {
  try {
    mfAssert(sp, "BasicParticleSystem: null solver");
    Solver* s = S(sp);
    // z-slab partitioned solvers: the particle system is REPLICATED -- every rank holds all particles and runs the particle-side
    // kernels (advection, grid -> particle transfers read the grids through the one address space), which are deterministic, so the
    // replicas stay bit-identical without any exchange; the particle -> grid writers (mapPartsToMAC, markFluidCells) run on rank 0
    // alone, whose stores reach the other ranks' slabs over NVLink, followed by the device barrier (flip.cpp:619-670 has no
    // counterpart: the reference is a shared-memory program).
    mfAssert(s->N < (1LL << 32) - 1, "BasicParticleSystem: grid too large for 32-bit face keys");
    Parts* p = new Parts();
    p->s = s; p->n = 0; p->cap = 0; p->pos = nullptr; p->flag = nullptr; p->scratch = nullptr; p->scratchBytes = 0;
    return reinterpret_cast<mf_parts*>(p);
  } catch (const std::exception& e) { set_last_error(e.what()); return nullptr; }
}
void mf_parts_destroy(mf_parts* pp) {
  if (!pp) return;
  Parts* p = P(pp);
  p->s->use();
  cudaStreamSynchronize(p->s->stream);
  cudaFree(p->pos); cudaFree(p->flag); cudaFree(p->scratch);
  for (PData& d : p->pdata) cudaFree(d.d);
  delete p;
}
long long mf_parts_size(const mf_parts* p) { return p ? P(p)->n : 0; }

int mf_parts_create_pdata(mf_parts* pp, int ncomp, int* index) {
  MF_API_BEGIN
  mfAssert(pp && index, "ParticleSystem::create: null argument");
  mfAssert(ncomp == 1 || ncomp == 3 || ncomp == -1, "ParticleSystem::create: unknown particle data type");
  Parts* p = P(pp); p->s->use();
  PData d; d.ncomp = ncomp; d.d = nullptr;
  const int nc = ncomp == 3 ? 3 : 1;
  if (p->cap) { CUDA_CHECK(cudaMalloc(&d.d, sizeof(float) * nc * (size_t)p->cap)); CUDA_CHECK(cudaMemset(d.d, 0, sizeof(float) * nc * (size_t)p->cap)); }
  p->pdata.push_back(d);
  *index = (int)p->pdata.size() - 1;
  MF_API_END
}

// replace the contents: pos = n x 3 (xyz interleaved) host floats, flag = n host ints (or NULL: PNEW); particle data is zeroed
// (ParticleDataImpl::initNewValue without a grid source, particle.cpp:349-356)
int mf_parts_upload(mf_parts* pp, const float* pos, const int* flag, long long n) {
  MF_API_BEGIN
  mfAssert(pp && (pos || n == 0) && n >= 0, "mf_parts_upload: bad argument");
  Parts* p = P(pp); Solver* s = p->s; s->use();
  p->n = 0;
  reserve(p, n);
  if (n) {
    std::vector<float> soa((size_t)3 * n);
    std::vector<int> fl((size_t)n, (int)PNEW);
    for (long long i = 0; i < n; i++) for (int c = 0; c < 3; c++) soa[(size_t)c * n + i] = pos[3 * i + c];
    if (flag) for (long long i = 0; i < n; i++) fl[i] = flag[i];
    for (int c = 0; c < 3; c++) CUDA_CHECK(cudaMemcpy(p->pos + c * p->cap, soa.data() + (size_t)c * n, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice));
    CUDA_CHECK(cudaMemcpy(p->flag, fl.data(), sizeof(int) * (size_t)n, cudaMemcpyHostToDevice));
    for (PData& d : p->pdata) CUDA_CHECK(cudaMemset(d.d, 0, sizeof(float) * (d.ncomp == 3 ? 3 : 1) * (size_t)p->cap));
  }
  p->n = n;
  MF_API_END
}
int mf_parts_download(const mf_parts* pp, float* pos, int* flag) {
  MF_API_BEGIN
  mfAssert(pp, "mf_parts_download: null particle system");
  const Parts* p = P(pp); Solver* s = p->s; s->use();
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  const long long n = p->n;
  if (!n) return 0;
  if (pos) {
    std::vector<float> soa((size_t)3 * n);
    for (int c = 0; c < 3; c++) CUDA_CHECK(cudaMemcpy(soa.data() + (size_t)c * n, p->pos + c * p->cap, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    for (long long i = 0; i < n; i++) for (int c = 0; c < 3; c++) pos[3 * i + c] = soa[(size_t)c * n + i];
  }
  if (flag) CUDA_CHECK(cudaMemcpy(flag, p->flag, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost));
  MF_API_END
}
// particle data <-> host: n x ncomp interleaved (float, or int for ncomp == -1)
int mf_pdata_download(const mf_parts* pp, int pd, void* out) {
  MF_API_BEGIN
  mfAssert(pp && out, "mf_pdata_download: null argument");
  const Parts* p = P(pp); Solver* s = p->s; s->use();
  mfAssert(pd >= 0 && pd < (int)p->pdata.size(), "mf_pdata_download: no such particle data");
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  const PData& d = p->pdata[pd];
  const long long n = p->n; const int nc = d.ncomp == 3 ? 3 : 1;
  if (!n) return 0;
  std::vector<float> soa((size_t)nc * n);
  for (int c = 0; c < nc; c++) CUDA_CHECK(cudaMemcpy(soa.data() + (size_t)c * n, d.d + c * p->cap, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
  float* o = static_cast<float*>(out);
  for (long long i = 0; i < n; i++) for (int c = 0; c < nc; c++) o[nc * i + c] = soa[(size_t)c * n + i];
  MF_API_END
}
int mf_pdata_upload(mf_parts* pp, int pd, const void* in) {
  MF_API_BEGIN
  mfAssert(pp && in, "mf_pdata_upload: null argument");
  Parts* p = P(pp); Solver* s = p->s; s->use();
  mfAssert(pd >= 0 && pd < (int)p->pdata.size(), "mf_pdata_upload: no such particle data");
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  PData& d = p->pdata[pd];
  const long long n = p->n; const int nc = d.ncomp == 3 ? 3 : 1;
  if (!n) return 0;
  std::vector<float> soa((size_t)nc * n);
  const float* i_ = static_cast<const float*>(in);
  for (long long i = 0; i < n; i++) for (int c = 0; c < nc; c++) soa[(size_t)c * n + i] = i_[nc * i + c];
  for (int c = 0; c < nc; c++) CUDA_CHECK(cudaMemcpy(d.d + c * p->cap, soa.data() + (size_t)c * n, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice));
  MF_API_END
}

// sampleFlagsWithParticles (flip.cpp:32-56): the jitter comes from ONE sequential random stream walked in cell order, so the
// positions are generated on the host (a set-up step) from the downloaded flags and appended to the system
int mf_sample_flags_with_particles(const mf_grid* flagsG, mf_parts* pp, int discretization, float randomness) {
  MF_API_BEGIN
  REQ_GRID(flagsG, G(flagsG)->isInt(), "flags", "sampleFlagsWithParticles");
  mfAssert(pp, "sampleFlagsWithParticles: argument 'parts' is missing");
  mfAssert(discretization > 0, "sampleFlagsWithParticles: discretization must be positive");
  Parts* p = P(pp); Solver* s = p->s; s->use();
  mfAssert(G(flagsG)->s == s, "sampleFlagsWithParticles: flags and parts belong to different solvers");
  std::vector<int> fl((size_t)s->N);
  check_api(mf_grid_download(flagsG, fl.data()));          // (every rank of a partitioned solver reads the whole grid and samples identically)
  const bool is3D = s->is3D;
  const float jlen = randomness / (float)discretization;
  const float disp = (float)(1.0 / (double)discretization);
  std::mt19937 gen(9832u);
  auto real = [&]() { return (float)((double)gen() * (1.0 / 4294967295.0)); };       // RandomStream::getReal (randomstream.h:385)
  std::vector<float> np;
  for (int k = 0; k < s->nz; k++)
    for (int j = 0; j < s->ny; j++)
      for (int i = 0; i < s->nx; i++) {
        const int f = fl[(size_t)i + (size_t)s->nx * j + (size_t)s->nx * s->ny * k];
        if ((f & MF_OBSTACLE) || !(f & MF_FLUID)) continue;
        for (int dk = 0; dk < (is3D ? discretization : 1); dk++)
          for (int dj = 0; dj < discretization; dj++)
            for (int di = 0; di < discretization; di++) {
              float q[3] = { (float)i + disp * (float)(0.5 + di), (float)j + disp * (float)(0.5 + dj), (float)k + disp * (float)(0.5 + dk) };
              const float a = real(), b = real(), c = real();               // getVec3(): x, y, z in this order
              q[0] += jlen * (1.f - (float)(2.0 * (double)a));
              q[1] += jlen * (1.f - (float)(2.0 * (double)b));
              q[2] += jlen * (1.f - (float)(2.0 * (double)c));
              if (!is3D) q[2] = 0.5f;
              np.push_back(q[0]); np.push_back(q[1]); np.push_back(q[2]);
            }
      }
  // insertBufferedParticles (particle.h:637-661): PNEW is cleared on the old particles, the new ones carry it
  const long long add = (long long)(np.size() / 3), old = p->n;
  std::vector<float> allPos((size_t)3 * (old + add));
  std::vector<int> allFlag((size_t)(old + add));
  if (old) { check_api(mf_parts_download(pp, allPos.data(), allFlag.data())); for (long long i = 0; i < old; i++) allFlag[i] &= ~(int)PNEW; }
  std::copy(np.begin(), np.end(), allPos.begin() + 3 * old);
  for (long long i = 0; i < add; i++) allFlag[old + i] = (int)PNEW;
  // keep the particle data of the old particles
  std::vector<std::vector<float>> keep(p->pdata.size());
  for (size_t d = 0; d < p->pdata.size(); d++) if (old) { keep[d].resize((size_t)(p->pdata[d].ncomp == 3 ? 3 : 1) * old); check_api(mf_pdata_download(pp, (int)d, keep[d].data())); }
  check_api(mf_parts_upload(pp, allPos.data(), allFlag.data(), old + add));
  for (size_t d = 0; d < p->pdata.size(); d++) if (old) {
    std::vector<float> full((size_t)(p->pdata[d].ncomp == 3 ? 3 : 1) * (old + add), 0.f);
    std::copy(keep[d].begin(), keep[d].end(), full.begin());
    check_api(mf_pdata_upload(pp, (int)d, full.data()));
  }
  MF_API_END
}

// ParticleSystem::advectInGrid (particle.h:527-550); ptype / exclude are not supported (NULL / 0 in the reference's scenes)
int mf_parts_advect_in_grid(mf_parts* pp, const mf_grid* flagsG, const mf_grid* velG, int integrationMode, int deleteInObstacle,
                            int stopInObstacle, int skipNew) {
  MF_API_BEGIN
  mfAssert(pp, "advectInGrid: null particle system");
  REQ_GRID(flagsG, G(flagsG)->isInt(), "flags", "advectInGrid"); REQ_GRID(velG, G(velG)->isMAC(), "vel", "advectInGrid");
  Parts* p = P(pp); Solver* s = p->s; s->use();
  mfAssert(G(flagsG)->s == s && G(velG)->s == s, "advectInGrid: grids and particles belong to different solvers");
  mfAssert(integrationMode >= 0 && integrationMode <= 2, "unknown integration type");
  const long long n = p->n;
  if (!n) return 0;
  const Geo g = geo(s);
  // scratch: x0 (3n), u (3n), total (3n)
  float* sc = (float*)scratchOf(p, sizeof(float) * 9 * (size_t)n);
  float* x0 = sc; float* u = sc + 3 * n; float* tot = sc + 6 * n;
  const int* fl = G(flagsG)->di(); const CV3 vel = cv3(G(velG));
  for (int c = 0; c < 3; c++) CUDA_CHECK(cudaMemcpyAsync(x0 + c * n, p->pos + c * p->cap, sizeof(float) * (size_t)n, cudaMemcpyDeviceToDevice, s->stream));
  // u must hold defined values where a not-stopped particle in an obstacle keeps "its previous value": the reference's return
  // vector starts value-initialised
  CUDA_CHECK(cudaMemsetAsync(u, 0, sizeof(float) * 3 * (size_t)n, s->stream));
  auto velocity = [&]() { MF_LAUNCH(s, k_part_velocity, pblocks(n), 256, g, n, (const float*)p->px(), (const float*)p->py(), (const float*)p->pz(), p->flag, vel, fl,
                                    s->dt, deleteInObstacle, stopInObstacle, skipNew, u, u + n, u + 2 * n); };
  auto stage = [&](float a, float* total, float b, int init) {
    for (int c = 0; c < 3; c++) MF_LAUNCH(s, k_part_stage, pblocks(n), 256, n, p->pos + c * p->cap, (const float*)(x0 + c * n), (const float*)(u + c * n), a,
                                          total ? total + c * n : (float*)nullptr, b, init); };
  velocity();
  if (integrationMode == 0) stage(1.f, nullptr, 0.f, 0);
  else if (integrationMode == 1) { stage(0.5f, nullptr, 0.f, 0); velocity(); stage(1.f, nullptr, 0.f, 0); }
  else {
    stage(0.5f, tot, 1.f, 1); velocity();
    stage(0.5f, tot, 2.f, 0); velocity();
    stage(1.f, tot, 2.f, 0); velocity();
    for (int c = 0; c < 3; c++) MF_LAUNCH(s, k_part_rk4_final, pblocks(n), 256, n, p->pos + c * p->cap, (const float*)(x0 + c * n), (const float*)(u + c * n), (const float*)(tot + c * n));
  }
  if (!deleteInObstacle) {
    MF_LAUNCH(s, k_part_clamp, pblocks(n), 256, g, n, p->px(), p->py(), p->pz(), (const int*)p->flag, fl, (const float*)x0, (const float*)(x0 + n), (const float*)(x0 + 2 * n), stopInObstacle);
  } else {
    MF_LAUNCH(s, k_part_delete_in_obstacle, pblocks(n), 256, g, n, (const float*)p->px(), (const float*)p->py(), (const float*)p->pz(), p->flag, fl);
  }
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_mark_fluid_cells(const mf_parts* pp, mf_grid* flagsG, const mf_grid* phiObs) {
  MF_API_BEGIN
  mfAssert(pp, "markFluidCells: argument 'parts' is missing");
  REQ_GRID(flagsG, G(flagsG)->isInt(), "flags", "markFluidCells");
  if (phiObs) REQ_GRID(phiObs, G(phiObs)->isReal(), "phiObs", "markFluidCells");
  const Parts* p = P(pp); Solver* s = p->s; s->use();
  Grid* fg = G(flagsG);
  mfAssert(fg->s == s, "markFluidCells: flags and parts belong to different solvers");
  markFlagsChanged(fg);
  const Geo g = geo(s);                       // particle kernels address the whole grid
  if (s->rankBegin == 0) {                    // the writer (see mf_parts_create)
    MF_LAUNCH(s, k_clear_fluid_flags, pblocks(s->N), 256, s->N, fg->di());
    if (p->n) MF_LAUNCH(s, k_mark_fluid, pblocks(p->n), 256, g, p->n, (const float*)p->px(), (const float*)p->py(), (const float*)p->pz(), (const int*)p->flag, fg->di());
  }
  s->rankSync();
  if (phiObs) {
    TmpReal tmp(s);
    int* nf = reinterpret_cast<int*>(tmp.p);
    MF_RANKS(s, r) { const Span sp = span(s, r); if (sp.n) MF_LAUNCH(s, k_copy_words, pblocks(sp.n), 256, sp.n, (unsigned*)(nf + sp.off), (const unsigned*)(fg->di() + sp.off)); }
    s->rankSync();
    MF_LAUNCH_IJK(s, k_set_nb_obstacle, nf, (const int*)fg->di(), (const float*)G(phiObs)->d);
    MF_RANKS(s, r) { const Span sp = span(s, r); if (sp.n) MF_LAUNCH(s, k_copy_words, pblocks(sp.n), 256, sp.n, (unsigned*)(fg->di() + sp.off), (const unsigned*)(nf + sp.off)); }
    s->rankSync();
  }
  MF_LAUNCH_CHECK();
  MF_API_END
}

static void gridToParts(const mf_grid* flagsG, const mf_grid* velG, const mf_grid* velOldG, const mf_parts* pp, int pdVel, int flip, float flipRatio, const char* fn) {
  mfAssert(pp, fn << ": argument 'parts' is missing");
  mfAssert(velG && G(velG)->isMAC(), fn << ": argument 'vel' has the wrong grid type or is missing");
  if (flip) mfAssert(velOldG && G(velOldG)->isMAC(), fn << ": argument 'velOld' has the wrong grid type or is missing");
  const Parts* p = P(pp); Solver* s = p->s; s->use();
  (void)flagsG;
  const PData& d = pdataOf(p, pdVel, 3, fn);
  if (!p->n) return;
  const Geo g = geo(s);
  const CV3 v = cv3(G(velG)); const CV3 vo = flip ? cv3(G(velOldG)) : v;
  MF_LAUNCH(s, k_grid_to_parts, pblocks(p->n), 256, g, p->n, (const float*)p->px(), (const float*)p->py(), (const float*)p->pz(), (const int*)p->flag, v, vo, flip, flipRatio,
            d.d, d.d + p->cap, d.d + 2 * p->cap);
}
int mf_map_mac_to_parts(const mf_grid* flags, const mf_grid* vel, const mf_parts* parts, int pdVel) {
  MF_API_BEGIN
  gridToParts(flags, vel, nullptr, parts, pdVel, 0, 0.f, "mapMACToParts");
  MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_flip_velocity_update(const mf_grid* flags, const mf_grid* vel, const mf_grid* velOld, const mf_parts* parts, int pdVel, float flipRatio) {
  MF_API_BEGIN
  gridToParts(flags, vel, velOld, parts, pdVel, 1, flipRatio, "flipVelocityUpdate");
  MF_LAUNCH_CHECK();
  MF_API_END
}

// mapPartsToMAC (flip.cpp:637-661)
int mf_map_parts_to_mac(const mf_grid* flagsG, mf_grid* velG, mf_grid* velOldG, const mf_parts* pp, int pdVel, mf_grid* weightG) {
  MF_API_BEGIN
  (void)flagsG;
  mfAssert(pp, "mapPartsToMAC: argument 'parts' is missing");
  REQ_GRID(velG, G(velG)->isMAC(), "vel", "mapPartsToMAC"); REQ_GRID(velOldG, G(velOldG)->isMAC(), "velOld", "mapPartsToMAC");
  if (weightG) REQ_GRID(weightG, G(weightG)->isVec(), "weight", "mapPartsToMAC");
  const Parts* p = P(pp); Solver* s = p->s; s->use();
  mfAssert(G(velG)->s == s && G(velOldG)->s == s, "mapPartsToMAC: grids and particles belong to different solvers");
  const PData& d = pdataOf(p, pdVel, 3, "mapPartsToMAC");
  TmpVec tmpW(s);
  float* weight = weightG ? G(weightG)->d : tmpW.p;
  float* vel = G(velG)->d;
  const bool writer = s->rankBegin == 0;      // see mf_parts_create: rank 0 scatters, its stores reach every slab; then the device barrier
  const long long n = p->n, m = 8 * n;
  if (writer) {
    MF_LAUNCH(s, k_zero_words, pblocks(3 * s->Np), 256, 3 * s->Np, (unsigned*)weight);
    MF_LAUNCH(s, k_zero_words, pblocks(3 * s->Np), 256, 3 * s->Np, (unsigned*)vel);
  }
  if (n && writer) {
    const Geo g = geo(s);
    size_t sortBytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sortBytes, (const unsigned*)nullptr, (unsigned*)nullptr, (const unsigned*)nullptr, (unsigned*)nullptr, (int)m, 0, 32, s->stream);
    const size_t arr = ((size_t)m * 4 + 255) / 256 * 256;
    char* sc = (char*)scratchOf(const_cast<Parts*>(p), 6 * arr + sortBytes + 256);
    unsigned* keys = (unsigned*)sc; unsigned* ids = (unsigned*)(sc + arr); unsigned* keys2 = (unsigned*)(sc + 2 * arr); unsigned* ids2 = (unsigned*)(sc + 3 * arr);
    float* w = (float*)(sc + 4 * arr); float* wv = (float*)(sc + 5 * arr); void* sortTmp = sc + 6 * arr;
    mfAssert(m < (1LL << 31), "mapPartsToMAC: too many particles for one pass");
    for (int c = 0; c < 3; c++) {
      const float* val = d.d + c * p->cap;
      if (c == 0) MF_LAUNCH(s, k_part_contrib<0>, pblocks(n), 256, g, n, (const float*)p->px(), (const float*)p->py(), (const float*)p->pz(), (const int*)p->flag, val, keys, ids, w, wv);
      else if (c == 1) MF_LAUNCH(s, k_part_contrib<1>, pblocks(n), 256, g, n, (const float*)p->px(), (const float*)p->py(), (const float*)p->pz(), (const int*)p->flag, val, keys, ids, w, wv);
      else MF_LAUNCH(s, k_part_contrib<2>, pblocks(n), 256, g, n, (const float*)p->px(), (const float*)p->py(), (const float*)p->pz(), (const int*)p->flag, val, keys, ids, w, wv);
      CUDA_CHECK(cub::DeviceRadixSort::SortPairs(sortTmp, sortBytes, (const unsigned*)keys, keys2, (const unsigned*)ids, ids2, (int)m, 0, 32, s->stream));
      MF_LAUNCH(s, k_part_accumulate, pblocks(m), 256, m, (const unsigned*)keys2, (const unsigned*)ids2, (const float*)w, (const float*)wv, weight + c * s->Np, vel + c * s->Np);
    }
  }
  if (writer) {
    MF_LAUNCH(s, k_stomp_divide, pblocks(3 * s->Np), 256, 3 * s->Np, vel, weight);
    MF_LAUNCH(s, k_copy_words, pblocks(3 * s->Np), 256, 3 * s->Np, (unsigned*)G(velOldG)->d, (const unsigned*)vel);
  }
  s->rankSync();
  MF_LAUNCH_CHECK();
  MF_API_END
}

} // extern "C"
