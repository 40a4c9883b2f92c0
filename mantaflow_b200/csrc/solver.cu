// solver.cu -- FluidSolver / Grid<T> data model on the device, elementwise ops and reductions.
// Reference: source/fluidsolver.{h,cpp} (temp-grid pool :25-50, ctor :107), source/grid.{h,cpp}.
#include "common.cuh"
#include <atomic>
#include "reduce.cuh"
#include "comm.h"
#include <cstring>
#include <cmath>
#include <algorithm>

namespace mf { struct Grid; struct Solver; }
void launch_norm_square(mf::Solver* s, const mf::Grid* g, float* o);

namespace mf {
unsigned long long nextGridVersion() { static unsigned long long v = 0; return ++v; }   // (API calls are single-threaded per process)

static thread_local std::string g_lastError;
void set_last_error(const std::string& s) { g_lastError = s; }

void check_same_solver(const Grid* a, const Grid* b, const char* what) {
  if (a && b && a->s != b->s) mfErr(what << ": grids belong to different solvers");
}

void Solver::use() const { cudaSetDevice(device); }

// ---- multi-GPU plumbing (comm.h) -------------------------------------------------------------------------
void* Solver::galloc(size_t bytesPerComp, int ncomp, size_t compStride, const long long* elemSplit, int elemSize) {
  if (comm && !comm_emulated(comm)) {
    size_t bs[COMM_MAX_RANKS + 1];
    for (int r = 0; r <= nranks; r++)
      bs[r] = (size_t)(elemSplit ? elemSplit[r] : (long long)nx * ny * zsplit[r]) * (size_t)elemSize;
    return comm_alloc(comm, bytesPerComp, ncomp, compStride, bs);
  }
  const size_t total = ncomp > 1 ? compStride * (size_t)(ncomp - 1) + bytesPerComp : bytesPerComp;
  void* p = nullptr;
  CUDA_CHECK(cudaMalloc(&p, total));
  CUDA_CHECK(cudaMemsetAsync(p, 0, total, stream));
  return p;
}
void Solver::gfree(void* p) {
  if (!p) return;
  if (comm && !comm_emulated(comm)) comm_free(comm, p); else cudaFree(p);
}
void Solver::rankSync() { if (comm && nranks > 1) comm_barrier(comm, stream); }
void Solver::clearShared(void* p, size_t bytes) {
  if (!multi() || comm_emulated(comm)) { CUDA_CHECK(cudaMemsetAsync(p, 0, bytes, stream)); return; }
  rankSync();                                   // nobody still reads the previous contents
  if (rankBegin == 0) CUDA_CHECK(cudaMemsetAsync(p, 0, bytes, stream));
  rankSync();
}
RedCtx redCtx(Solver* s, int rank) {
  RedCtx c; c.partials = s->redScratch; c.counter = s->redCounter; c.nranks = s->nranks; c.rank = rank;
  c.shared = nullptr; c.stride = 0; c.parity = s->redParity; c.self = s->rankBegin;
  if (s->multi()) {
    c.shared = reinterpret_cast<char*>(comm_shared(s->comm, 0));
    c.stride = (size_t)(reinterpret_cast<char*>(comm_shared(s->comm, 1)) - c.shared);
  }
  return c;
}
static_assert(COMM_MAX_RANKS == 16 && COMM_RED_K == 8 && offsetof(CommShared, red) == 16 * sizeof(unsigned long long),
              "reduce.cuh redSlot() hard-codes the CommShared layout");

float* Solver::getReal() {
  if (!poolReal.empty()) { float* p = poolReal.back(); poolReal.pop_back(); return p; }
  float* p = (float*)galloc(sizeof(float) * (size_t)(N + 64));
  poolBytes += sizeof(float) * (N + 64);
  return p;
}
void Solver::putReal(float* p) { poolReal.push_back(p); }
float* Solver::getVec() {
  if (!poolVec.empty()) { float* p = poolVec.back(); poolVec.pop_back(); return p; }
  float* p = (float*)galloc(sizeof(float) * (size_t)Np, 3, sizeof(float) * (size_t)Np);
  poolBytes += sizeof(float) * (3 * Np);
  return p;
}
void Solver::putVec(float* p) { poolVec.push_back(p); }
cudaEvent_t Solver::profEvent() {
  if (!evPool.empty()) { cudaEvent_t e = evPool.back(); evPool.pop_back(); return e; }
  cudaEvent_t e; CUDA_CHECK(cudaEventCreate(&e)); return e;
}

struct SmallPayload { unsigned w[16]; };
// payload into mapped pinned memory, then (after a system-scope fence) the ticket the host is spinning on
__global__ void k_read_small(unsigned* dst, const unsigned* __restrict__ src, int nwords, volatile unsigned* seq, unsigned ticket) {
  if ((int)threadIdx.x < nwords) dst[threadIdx.x] = src[threadIdx.x];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) *seq = ticket;
}
__global__ void k_write_small(unsigned* dst, SmallPayload p, int nwords) {
  if ((int)threadIdx.x < nwords) dst[threadIdx.x] = p.w[threadIdx.x];
}
void Solver::readSmall(void* pinnedDst, const void* dev, int bytes) { readSmallWait(readSmallLaunch(pinnedDst, dev, bytes)); }
// The two halves of readSmall: the copy kernel is queued now, the host waits for its ticket later -- kernels launched in between
// run behind the copy, so the host can decide about the work that follows while the device is already busy with its first part.
unsigned Solver::readSmallLaunch(void* pinnedDst, const void* dev, int bytes) {
  void* mapped = nullptr;
  CUDA_CHECK(cudaHostGetDevicePointer(&mapped, pinnedDst, 0));
  // The host waits on a ticket in pinned memory instead of cudaStreamSynchronize (the PCG reads its stop flag once per iteration:
  // the wake-up latency of a stream synchronisation is paid 4-5 times per step); every 64 Ki polls the stream is queried so that a
  // faulted kernel surfaces as an error instead of a hang.  MF_SYNC_READS=1 restores the plain synchronisation.
  void* seqDev = nullptr;
  CUDA_CHECK(cudaHostGetDevicePointer(&seqDev, (void*)seqHost, 0));
  const unsigned ticket = ++seqTicket;
  k_read_small<<<1, 64, 0, stream>>>((unsigned*)mapped, (const unsigned*)dev, bytes / 4, (volatile unsigned*)seqDev, ticket);
  launches++;
  CUDA_CHECK(cudaGetLastError());
  return ticket;
}
void Solver::readSmallWait(unsigned ticket) {
  static const bool plain = getenv("MF_SYNC_READS") != nullptr;
  if (plain) { CUDA_CHECK(cudaStreamSynchronize(stream)); return; }
  volatile unsigned* seq = seqHost;
  for (unsigned spins = 1; *seq != ticket; spins++) {
    if ((spins & 0xffffu) == 0) {
      const cudaError_t e = cudaStreamQuery(stream);
      if (e == cudaSuccess) break;                                   // everything queued has finished: the ticket is written
      if (e != cudaErrorNotReady) CUDA_CHECK(e);
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
}
void Solver::writeSmall(void* dev, const void* hostSrc, int bytes) {
  SmallPayload p; memset(&p, 0, sizeof(p)); memcpy(&p, hostSrc, (size_t)bytes);
  k_write_small<<<1, 32, 0, stream>>>((unsigned*)dev, p, bytes / 4);
  launches++;
}
void fetchRed(Solver* s, int count) { s->readSmall(s->redHost, s->redOut, (int)sizeof(double) * count); }

// ---- elementwise kernels ---------------------------------------------------------------
__global__ void k_fill_f(float* d, float v, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) d[i] = v;
}
__global__ void k_fill_i(int* d, int v, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) d[i] = v;
}
// me += f * other   (gridScaledAdd, grid.h:514)
__global__ void k_scaled_add(float* me, const float* other, float f, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) me[i] += f * other[i];
}
__global__ void k_mult_const(float* me, float f, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) me[i] *= f;
}
__global__ void k_aos_to_soa(const float* __restrict__ aos, float* x, float* y, float* z, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) { x[i] = aos[3 * i]; y[i] = aos[3 * i + 1]; z[i] = aos[3 * i + 2]; }
}
__global__ void k_soa_to_aos(float* __restrict__ aos, const float* x, const float* y, const float* z, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) { aos[3 * i] = x[i]; aos[3 * i + 1] = y[i]; aos[3 * i + 2] = z[i]; }
}

static inline int ewBlocks(const Solver* s, long long n) {
  long long b = (n + 255) / 256;
  long long cap = (long long)s->numSMs * 16;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

// ---- reductions --------------------------------------------------------------------------
// what: 0 dot (float product, double accumulate; conjugategrad.cpp:177), 1 sum of squares in double, 2 min/max
template <int WHAT, class T>
__global__ void k_reduce(const T* __restrict__ a, const T* __restrict__ b, long long n, RedCtx ctx, double* out) {
  __shared__ double sh[32];
  double acc[2] = { WHAT == 2 ? 1e300 : 0.0, -1e300 };
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    if (WHAT == 0) acc[0] += (double)((float)a[i] * (float)b[i]);
    else if (WHAT == 1) { const double v = (double)a[i]; acc[0] += v * v; }
    else { const double v = (double)a[i]; if (v < acc[0]) acc[0] = v; if (v > acc[1]) acc[1] = v; }
  }
  const int ops[2] = { WHAT == 2 ? 1 : 0, 2 };
  gridReduceK<2>(acc, ops, ctx, sh, FinStore<2>{ out });
}
// max over cells of |a-b| (Real, gridMaxDiff grid.cpp:437) or sum of component |diffs| (Vec3, :457)
template <class T>
__global__ void k_maxdiff(const T* __restrict__ a, const T* __restrict__ b, long long n, long long np, int ncomp, RedCtx ctx, double* out) {
  __shared__ double sh[32];
  double mx = 0.0;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    double d = 0.0;
    for (int c = 0; c < ncomp; c++) d += fabs((double)a[i + c * np] - (double)b[i + c * np]);
    if (d > mx) mx = d;
  }
  double acc[1] = { mx };
  const int ops[1] = { 2 };
  gridReduceK<1>(acc, ops, ctx, sh, FinStore<1>{ out });
}

template <int WHAT, class T>
static void reduce2(Solver* s, const T* a, const T* b, int nfetch) {
  MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH(s, (k_reduce<WHAT, T>), redBlocks(s, sp.n), RED_THREADS, a + sp.off, b ? b + sp.off : nullptr, sp.n, redCtx(s, r), s->redOut); }
  redFinish<2>(s, RedOps<2>{ { WHAT == 2 ? 1 : 0, 2 } }, FinStore<2>{ s->redOut });
  fetchRed(s, nfetch);
}
double reduce_dot(Solver* s, const float* a, const float* b, long long) { reduce2<0, float>(s, a, b, 1); return s->redHost[0]; }
double reduce_sumsqr(Solver* s, const float* a, long long) { reduce2<1, float>(s, a, nullptr, 1); return s->redHost[0]; }
void reduce_minmax(Solver* s, const float* a, long long, float* mn, float* mx) { reduce2<2, float>(s, a, nullptr, 2); *mn = (float)s->redHost[0]; *mx = (float)s->redHost[1]; }
void reduce_minmax_int(Solver* s, const int* a, long long, int* mn, int* mx) { reduce2<2, int>(s, a, nullptr, 2); *mn = (int)s->redHost[0]; *mx = (int)s->redHost[1]; }

// elementwise kernel over every local rank's slab of one array (or of each listed array pair)
#define MF_EW(s, kern, ...) do { MF_RANKS(s, _r) { const Span sp = span(s, _r); if (sp.n) { MF_LAUNCH(s, kern, ewBlocks(s, sp.n), 256, __VA_ARGS__); } } } while (0)

// per-rank slab of every component: f(component base pointer offset to the slab, cells in the slab)
template <class F> static void forSlabs(Solver* s, const Grid* g, F f) {
  const int nc = g->isVec() ? 3 : 1;
  MF_RANKS(s, r) { const Span sp = span(s, r); if (sp.n) for (int c = 0; c < nc; c++) f((g->isVec() ? g->comp(c) : g->d) + sp.off, sp.n, c, sp); }
  s->rankSync();
}

} // namespace mf

using namespace mf;

extern "C" {

int mf_version(void) { return 100; }
const char* mf_last_error(void) { return g_lastError.c_str(); }
int mf_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n; }
const char* mf_build_info(void) { return "mantaflow-b200 0.1 linux 64bit fp1 cuda sm_100a"; }

static Solver* createSolver(int nx, int ny, int nz, int dim, int device, Comm* comm) {
  mfAssert(dim == 2 || dim == 3, "FluidSolver: dim must be 2 or 3");
  mfAssert(nx > 0 && ny > 0 && nz > 0, "FluidSolver: invalid grid size");
  if (dim == 2) nz = 1;   // fluidsolver.cpp:107-112 forces z=1 in 2-D
  int ndev = mf_device_count();
  mfAssert(ndev > 0, "mantacuda: no CUDA device available (there is no CPU fallback)");
  mfAssert(device >= 0 && device < ndev, "mantacuda: invalid device index " << device);
  CUDA_CHECK(cudaSetDevice(device));
  Solver* s = new Solver();
  s->nx = nx; s->ny = ny; s->nz = nz; s->dim = dim; s->is3D = (dim == 3);
  s->N = (long long)nx * ny * nz;
  s->comm = comm; s->nranks = comm ? comm_nranks(comm) : 1; s->redParity = 0;
  if (comm && comm_emulated(comm)) { s->rankBegin = 0; s->rankEnd = s->nranks; }
  else { s->rankBegin = comm ? comm_rank(comm) : 0; s->rankEnd = s->rankBegin + 1; }
  mfAssert(s->nranks == 1 || (s->is3D && nz >= 2 * s->nranks), "FluidSolver: a grid partitioned over " << s->nranks << " GPUs needs dim == 3 and at least 2 z planes per GPU");
  for (int r = 0; r <= s->nranks; r++) s->zsplit[r] = (int)((long long)nz * r / s->nranks);
  // SoA component stride: 64 floats of slack behind every component; a whole number of allocation granules when the
  // components are mapped piecewise over several GPUs
  const long long unit = (comm && !comm_emulated(comm)) ? (long long)(comm_granularity(comm) / sizeof(float)) : 32;
  s->Np = (s->N + 64 + unit - 1) / unit * unit;
  s->strideZ = s->is3D ? (long long)nx * ny : 0;
  s->dt = 1.0f;
  int mx = nx > ny ? nx : ny; if (nz > mx) mx = nz;
  s->dx = (float)(1.0 / mx);       // grid.cpp:57
  s->device = device;
  s->profTag = -1; s->profLevels = getenv("MF_PROFILE_LEVELS") != nullptr;
  s->launches = 0; s->poolBytes = 0; s->aosStage = nullptr; s->mg = nullptr; s->mic = nullptr; s->cgCode = nullptr; s->codeFlags = nullptr; s->codeVersion = 0; s->codeFixing = 0; s->codeFixedCell = -1; s->noiseTile = nullptr; s->profiling = false; s->copyIn = nullptr; s->copyOut = nullptr;
  for (int q = 0; q < 16; q++) s->userEvents[q] = nullptr;
  CUDA_CHECK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  cudaDeviceProp prop; CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
  s->numSMs = prop.multiProcessorCount;
  CUDA_CHECK(cudaMalloc(&s->redScratch, sizeof(double) * 8 * RED_MAX_BLOCKS));
  CUDA_CHECK(cudaMalloc(&s->redCounter, sizeof(unsigned int) * 4));
  CUDA_CHECK(cudaMemset(s->redCounter, 0, sizeof(unsigned int) * 4));
  CUDA_CHECK(cudaMalloc(&s->redOut, sizeof(double) * 16));
  CUDA_CHECK(cudaMemset(s->redOut, 0, sizeof(double) * 16));
  CUDA_CHECK(cudaHostAlloc(&s->redHost, sizeof(double) * 16, cudaHostAllocMapped));
  CUDA_CHECK(cudaMalloc(&s->localInts, sizeof(int) * 16));
  CUDA_CHECK(cudaMemset(s->localInts, 0, sizeof(int) * 16));
  s->devFlagsScratch = (int*)s->galloc(sizeof(int) * 16);      // job-wide: kernels on any rank raise flags here
  CUDA_CHECK(cudaHostAlloc(&s->hostInts, sizeof(int) * 16, cudaHostAllocMapped));
  CUDA_CHECK(cudaHostAlloc((void**)&s->seqHost, 64, cudaHostAllocMapped)); *s->seqHost = 0; s->seqTicket = 0;
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  return s;
}

mf_solver* mf_solver_create(int nx, int ny, int nz, int dim, int device) {
  try { return reinterpret_cast<mf_solver*>(createSolver(nx, ny, nz, dim, device, nullptr)); }
  catch (const std::exception& e) { set_last_error(e.what()); return nullptr; }
}
// z-slab partitioned solver: every rank of the communicator makes the same call (and every later call on the solver
// and its grids, in the same order)
mf_solver* mf_solver_create_dist(mf_comm* comm, int nx, int ny, int nz, int dim) {
  try {
    mfAssert(comm, "mf_solver_create_dist: null communicator");
    Comm* c = reinterpret_cast<Comm*>(comm);
    return reinterpret_cast<mf_solver*>(createSolver(nx, ny, nz, dim, comm_device(c), c));
  } catch (const std::exception& e) { set_last_error(e.what()); return nullptr; }
}
int mf_solver_slab(const mf_solver* sp, int rank, int* k0, int* k1) {
  MF_API_BEGIN
  mfAssert(sp && k0 && k1, "mf_solver_slab: null argument");
  const Solver* s = S(sp);
  if (rank < 0) rank = s->rankBegin;
  mfAssert(rank < s->nranks, "mf_solver_slab: no such rank");
  *k0 = s->zsplit[rank]; *k1 = s->zsplit[rank + 1];
  MF_API_END
}
int mf_solver_ranks(const mf_solver* sp) { return sp ? S(sp)->nranks : 0; }
long long mf_solver_barrier_count(const mf_solver* sp) { return (sp && S(sp)->comm) ? comm_barriers_issued(S(sp)->comm) : 0; }

void mf_solver_destroy(mf_solver* sp) {
  if (!sp) return;
  Solver* s = S(sp);
  s->use();
  cudaStreamSynchronize(s->stream);
  release_mg(s);
  release_mic(s);
  for (float* p : s->poolReal) s->gfree(p);
  for (float* p : s->poolVec) s->gfree(p);
  if (s->cgCode) s->gfree(s->cgCode);
  if (s->noiseTile) cudaFree(s->noiseTile);
  cudaFree(s->redScratch); cudaFree(s->redCounter); cudaFree(s->redOut); cudaFreeHost(s->redHost);
  cudaFree(s->localInts); s->gfree(s->devFlagsScratch); cudaFreeHost(s->hostInts); cudaFreeHost((void*)s->seqHost);
  if (s->aosStage) cudaFree(s->aosStage);
  for (auto& r : s->profRecs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  for (cudaEvent_t e : s->evPool) cudaEventDestroy(e);
  for (int q = 0; q < 16; q++) if (s->userEvents[q]) cudaEventDestroy(s->userEvents[q]);
  if (s->copyIn) cudaStreamDestroy(s->copyIn);
  if (s->copyOut) cudaStreamDestroy(s->copyOut);
  cudaStreamDestroy(s->stream);
  delete s;
}

int mf_solver_set_dt(mf_solver* s, float dt) { MF_API_BEGIN mfAssert(s, "null solver"); S(s)->dt = dt; MF_API_END }
float mf_solver_get_dt(const mf_solver* s) { return s ? S(s)->dt : 0.f; }
int mf_solver_sync(mf_solver* s) {
  MF_API_BEGIN
  mfAssert(s, "null solver"); S(s)->use();
  CUDA_CHECK(cudaStreamSynchronize(S(s)->stream)); CUDA_CHECK(cudaGetLastError());
  mfAssert(!(S(s)->comm && comm_error(S(s)->comm)), "mantacuda: a cross-GPU barrier timed out (a peer rank died or made a different sequence of calls)");
  MF_API_END
}
void* mf_solver_stream(mf_solver* s) { return s ? (void*)S(s)->stream : nullptr; }
long long mf_solver_launch_count(const mf_solver* s) { return s ? S(s)->launches : 0; }
long long mf_solver_pool_bytes(const mf_solver* s) { return s ? S(s)->poolBytes : 0; }

int mf_solver_profile_begin(mf_solver* sp) {
  MF_API_BEGIN
  mfAssert(sp, "null solver");
  Solver* s = S(sp); s->use();
  for (auto& r : s->profRecs) { s->evPool.push_back(r.a); s->evPool.push_back(r.b); }
  s->profRecs.clear();
  s->profiling = true;
  MF_API_END
}
// stops recording and writes one line per kernel, "name launches total_ms\n", sorted by total time
int mf_solver_profile_end(mf_solver* sp, char* buf, int buflen) {
  MF_API_BEGIN
  mfAssert(sp && buf && buflen > 0, "mf_solver_profile_end: null argument");
  Solver* s = S(sp); s->use();
  s->profiling = false;
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  std::map<std::string, std::pair<long long, double>> agg;
  for (auto& r : s->profRecs) {
    float ms = 0.f; CUDA_CHECK(cudaEventElapsedTime(&ms, r.a, r.b));
    std::string n(r.name);
    // "(k_sl<NC, OS, OT>)" -> "k_sl"
    size_t b0 = n.find_first_not_of("( "); if (b0 == std::string::npos) b0 = 0;
    size_t e0 = n.find_first_of("<)", b0); n = n.substr(b0, e0 == std::string::npos ? std::string::npos : e0 - b0);
    if (n.compare(0, 4, "mf::") == 0) n = n.substr(4);
    if (r.tag >= 0 && s->profLevels) n += "@L" + std::to_string(r.tag);
    auto& a = agg[n]; a.first++; a.second += ms;
    s->evPool.push_back(r.a); s->evPool.push_back(r.b);
  }
  s->profRecs.clear();
  std::vector<std::pair<double, std::string>> order;
  for (auto& kv : agg) order.push_back({ -kv.second.second, kv.first });
  std::sort(order.begin(), order.end());
  std::ostringstream o;
  for (auto& e : order) { auto& a = agg[e.second]; o << e.second << " " << a.first << " " << a.second << "\n"; }
  std::string out = o.str();
  if ((int)out.size() >= buflen) out.resize(buflen - 1);
  memcpy(buf, out.c_str(), out.size() + 1);
  MF_API_END
}

// CUDA-event stopwatch on the solver stream (slots 0..15): callers time whole steps without touching the CUDA runtime
int mf_solver_event_record(mf_solver* sp, int slot) {
  MF_API_BEGIN
  mfAssert(sp && slot >= 0 && slot < 16, "mf_solver_event_record: bad argument");
  Solver* s = S(sp); s->use();
  if (!s->userEvents[slot]) CUDA_CHECK(cudaEventCreate(&s->userEvents[slot]));
  CUDA_CHECK(cudaEventRecord(s->userEvents[slot], s->stream));
  MF_API_END
}
int mf_solver_event_elapsed_ms(mf_solver* sp, int from, int to, float* ms) {
  MF_API_BEGIN
  mfAssert(sp && ms && from >= 0 && from < 16 && to >= 0 && to < 16, "mf_solver_event_elapsed_ms: bad argument");
  Solver* s = S(sp); s->use();
  mfAssert(s->userEvents[from] && s->userEvents[to], "mf_solver_event_elapsed_ms: event not recorded");
  CUDA_CHECK(cudaEventSynchronize(s->userEvents[to]));
  CUDA_CHECK(cudaEventElapsedTime(ms, s->userEvents[from], s->userEvents[to]));
  MF_API_END
}

void* mf_host_alloc(long long bytes) {
  void* p = nullptr;
  if (bytes <= 0 || cudaMallocHost(&p, (size_t)bytes) != cudaSuccess) { cudaGetLastError(); set_last_error("mf_host_alloc: cudaMallocHost failed"); return nullptr; }
  return p;
}
void mf_host_free(void* p) { if (p) cudaFreeHost(p); }

mf_grid* mf_grid_create(mf_solver* sp, int type) {
  try {
    mfAssert(sp, "Grid: null solver");
    Solver* s = S(sp); s->use();
    Grid* g = new Grid();
    g->s = s; g->type = type; g->pooled = false; g->d = nullptr; markFlagsChanged(g); g->xfer = nullptr;
    // Grid ctor clears (grid.cpp:58): galloc zero-fills
    if (type & MF_TYPE_VEC3) g->d = (float*)s->galloc(sizeof(float) * (size_t)s->Np, 3, sizeof(float) * (size_t)s->Np);
    else g->d = (float*)s->galloc(sizeof(float) * (size_t)(s->N + 64));
    return reinterpret_cast<mf_grid*>(g);
  } catch (const std::exception& e) { set_last_error(e.what()); return nullptr; }
}
}  // extern "C"
namespace mf { void freeXfer(Grid* g); }
extern "C" {
void mf_grid_destroy(mf_grid* gp) {
  if (!gp) return;
  Grid* g = G(gp);
  g->s->use();
  cudaStreamSynchronize(g->s->stream);
  freeXfer(g);
  g->s->gfree(g->d);
  delete g;
}
int mf_grid_type(const mf_grid* g) { return g ? G(g)->type : 0; }

int mf_grid_clear(mf_grid* gp) {
  MF_API_BEGIN
  mfAssert(gp, "Grid::clear: null grid");
  Grid* g = G(gp); Solver* s = g->s; s->use(); markFlagsChanged(g);
  forSlabs(s, g, [&](float* p, long long n, int, const Span&) { CUDA_CHECK(cudaMemsetAsync(p, 0, sizeof(float) * n, s->stream)); });
  MF_API_END
}
int mf_grid_copy(mf_grid* dp, const mf_grid* sp_) {
  MF_API_BEGIN
  mfAssert(dp && sp_, "Grid::copyFrom: null grid");
  Grid* d = G(dp); const Grid* a = G(sp_); Solver* s = d->s; s->use();
  check_same_solver(d, a, "Grid::copyFrom");
  mfAssert(d->isVec() == a->isVec() && d->isInt() == a->isInt(), "Grid::copyFrom: grid types do not match");
  forSlabs(s, d, [&](float* p, long long n, int c, const Span& sp) {
    const float* src = (a->isVec() ? a->comp(c) : a->d) + sp.off;
    CUDA_CHECK(cudaMemcpyAsync(p, src, sizeof(float) * n, cudaMemcpyDeviceToDevice, s->stream)); });
  { const int o = a->outflowState; markFlagsChanged(d); d->outflowState = o; }
  MF_API_END
}
// getComponent / setComponent (grid.cpp:743-751): one plane of the SoA vector grid <-> a real grid
static void componentCopy(const Grid* vec, const Grid* real, int component, bool toVec, const char* fn) {
  mfAssert(vec && real && vec->isVec() && real->isReal(), fn << ": needs a Vec3 grid and a Real grid");
  mfAssert(component >= 0 && component < 3, fn << ": component out of range");
  check_same_solver(vec, real, fn);
  Solver* s = vec->s; s->use();
  MF_RANKS(s, r) {
    const Span sp = span(s, r);
    if (!sp.n) continue;
    float* v = vec->comp(component) + sp.off; float* q = real->d + sp.off;
    CUDA_CHECK(cudaMemcpyAsync(toVec ? v : q, toVec ? q : v, sizeof(float) * sp.n, cudaMemcpyDeviceToDevice, s->stream));
  }
  s->rankSync();
}
int mf_grid_get_component(const mf_grid* source, mf_grid* target, int component) {
  MF_API_BEGIN
  componentCopy(G(source), G(target), component, false, "getComponent");
  MF_API_END
}
int mf_grid_set_component(const mf_grid* source, mf_grid* target, int component) {
  MF_API_BEGIN
  componentCopy(G(target), G(source), component, true, "setComponent");
  MF_API_END
}
int mf_grid_swap(mf_grid* ap, mf_grid* bp) {
  MF_API_BEGIN
  mfAssert(ap && bp, "Grid::swap: null grid");
  Grid* a = G(ap); Grid* b = G(bp);
  check_same_solver(a, b, "Grid::swap");
  mfAssert(a->isVec() == b->isVec() && a->isInt() == b->isInt(), "Grid::swap(): Grid types mismatch.");
  float* t = a->d; a->d = b->d; b->d = t;
  { const int oa = a->outflowState, ob = b->outflowState; markFlagsChanged(a); markFlagsChanged(b); a->outflowState = ob; b->outflowState = oa; }
  MF_API_END
}
// host <-> device.  `host` is the FULL grid in the reference layout on every rank (slabOnly == 0: each rank uploads its own
// slab from it; a download returns the whole grid to every rank, peers' slabs travelling over NVLink), or only the rank's own
// z-slab [k0,k1) (slabOnly == 1; mf_solver_slab).
static void gridUpload(Grid* g, const void* host, bool slabOnly) {
  Solver* s = g->s; s->use();
  markFlagsChanged(g);
  MF_RANKS(s, r) {
    const Span sp = span(s, r);
    if (!sp.n) continue;
    const long long hoff = slabOnly ? sp.off - span(s, s->rankBegin).off : sp.off;
    if (!g->isVec()) {
      CUDA_CHECK(cudaMemcpyAsync(g->d + sp.off, (const float*)host + hoff, sizeof(float) * (size_t)sp.n, cudaMemcpyHostToDevice, s->stream));
    } else {
      if (!s->aosStage) CUDA_CHECK(cudaMalloc(&s->aosStage, sizeof(float) * (size_t)(3 * span(s, s->rankBegin).n + 3 * (long long)s->nx * s->ny)));
      CUDA_CHECK(cudaMemcpyAsync(s->aosStage, (const float*)host + 3 * hoff, sizeof(float) * (size_t)(3 * sp.n), cudaMemcpyHostToDevice, s->stream));
      MF_LAUNCH(s, k_aos_to_soa, ewBlocks(s, sp.n), 256, s->aosStage, g->comp(0) + sp.off, g->comp(1) + sp.off, g->comp(2) + sp.off, sp.n);
      MF_LAUNCH_CHECK();
    }
  }
  s->rankSync();
  CUDA_CHECK(cudaStreamSynchronize(s->stream));  // host buffer may be reused by the caller
}
static void gridDownload(const Grid* g, void* host, bool slabOnly) {
  Solver* s = g->s; s->use();
  // all ranks have passed the barrier behind the last launch group, so every slab is complete
  const bool whole = !slabOnly && s->multi() && !comm_emulated(s->comm);
  for (int r = whole ? 0 : s->rankBegin; r < (whole ? s->nranks : s->rankEnd); r++) {
    const Span sp = span(s, r);
    if (!sp.n) continue;
    const long long hoff = slabOnly ? sp.off - span(s, s->rankBegin).off : sp.off;
    if (!g->isVec()) {
      CUDA_CHECK(cudaMemcpyAsync((float*)host + hoff, g->d + sp.off, sizeof(float) * (size_t)sp.n, cudaMemcpyDeviceToHost, s->stream));
    } else {
      if (!s->aosStage) CUDA_CHECK(cudaMalloc(&s->aosStage, sizeof(float) * (size_t)(3 * span(s, s->rankBegin).n + 3 * (long long)s->nx * s->ny)));
      MF_LAUNCH(s, k_soa_to_aos, ewBlocks(s, sp.n), 256, s->aosStage, g->comp(0) + sp.off, g->comp(1) + sp.off, g->comp(2) + sp.off, sp.n);
      MF_LAUNCH_CHECK();
      CUDA_CHECK(cudaMemcpyAsync((float*)host + 3 * hoff, s->aosStage, sizeof(float) * (size_t)(3 * sp.n), cudaMemcpyDeviceToHost, s->stream));
    }
  }
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  CUDA_CHECK(cudaGetLastError());
  s->rankSync();                                 // peers may overwrite the grid only after everybody has read it
}
// ---- streaming transfers: double-buffered staging, copies on their own streams ------------------------------------------
// in[k] / out[k] hold the process' slab in host layout.  Events order the three streams: the copy-in stream refills a slot only
// after the solver stream has consumed it (inFree), the solver stream overwrites an out slot only after the copy-out stream
// has drained it (outFree).
}  // extern "C"
namespace mf {
struct GridXfer {
  float* in[2]; float* out[2];
  cudaEvent_t inReady[2], inFree[2], outReady[2], outFree[2];
  bool inFreeSet[2], outFreeSet[2];
  int inHead, inCount, outNext;     // staged uploads form a queue of at most 2
  size_t bytes;
};
static Span localSpan(const Solver* s) {
  Span a = span(s, s->rankBegin), b = span(s, s->rankEnd - 1);
  Span r; r.off = a.off; r.n = b.off + b.n - a.off; return r;
}
static GridXfer* xferOf(const Grid* g) {
  Solver* s = g->s;
  if (!s->copyIn) { CUDA_CHECK(cudaStreamCreateWithFlags(&s->copyIn, cudaStreamNonBlocking)); CUDA_CHECK(cudaStreamCreateWithFlags(&s->copyOut, cudaStreamNonBlocking)); }
  if (g->xfer) return g->xfer;
  GridXfer* x = new GridXfer();
  memset(x, 0, sizeof(*x));
  x->bytes = sizeof(float) * (size_t)localSpan(s).n * (g->isVec() ? 3 : 1);
  for (int k = 0; k < 2; k++) {
    CUDA_CHECK(cudaEventCreateWithFlags(&x->inReady[k], cudaEventDisableTiming)); CUDA_CHECK(cudaEventCreateWithFlags(&x->inFree[k], cudaEventDisableTiming));
    CUDA_CHECK(cudaEventCreateWithFlags(&x->outReady[k], cudaEventDisableTiming)); CUDA_CHECK(cudaEventCreateWithFlags(&x->outFree[k], cudaEventDisableTiming));
  }
  g->xfer = x;
  return x;
}
void freeXfer(Grid* g) {
  GridXfer* x = g->xfer;
  if (!x) return;
  if (g->s->copyIn) { cudaStreamSynchronize(g->s->copyIn); cudaStreamSynchronize(g->s->copyOut); }
  for (int k = 0; k < 2; k++) {
    if (x->in[k]) cudaFree(x->in[k]);
    if (x->out[k]) cudaFree(x->out[k]);
    cudaEventDestroy(x->inReady[k]); cudaEventDestroy(x->inFree[k]); cudaEventDestroy(x->outReady[k]); cudaEventDestroy(x->outFree[k]);
  }
  delete x; g->xfer = nullptr;
}
}  // namespace mf
extern "C" {
int mf_grid_stage_upload(mf_grid* gp, const void* host) {
  MF_API_BEGIN
  mfAssert(gp && host, "Grid stage_upload: null argument");
  Grid* g = G(gp); Solver* s = g->s; s->use();
  GridXfer* x = xferOf(g);
  mfAssert(x->inCount < 2, "Grid stage_upload: two uploads are already staged -- commit one first");
  const int k = (x->inHead + x->inCount) & 1;
  if (!x->in[k]) CUDA_CHECK(cudaMalloc(&x->in[k], x->bytes));
  if (x->inFreeSet[k]) CUDA_CHECK(cudaStreamWaitEvent(s->copyIn, x->inFree[k], 0));
  CUDA_CHECK(cudaMemcpyAsync(x->in[k], host, x->bytes, cudaMemcpyHostToDevice, s->copyIn));
  CUDA_CHECK(cudaEventRecord(x->inReady[k], s->copyIn));
  x->inCount++;
  MF_API_END
}
int mf_grid_commit_upload(mf_grid* gp) {
  MF_API_BEGIN
  mfAssert(gp, "Grid commit_upload: null argument");
  Grid* g = G(gp); Solver* s = g->s; s->use();
  GridXfer* x = xferOf(g);
  mfAssert(x->inCount > 0, "Grid commit_upload: nothing staged");
  const int k = x->inHead;
  markFlagsChanged(g);
  CUDA_CHECK(cudaStreamWaitEvent(s->stream, x->inReady[k], 0));
  const long long base = localSpan(s).off;
  MF_RANKS(s, r) {
    const Span sp = span(s, r);
    if (!sp.n) continue;
    const long long hoff = sp.off - base;
    if (!g->isVec()) CUDA_CHECK(cudaMemcpyAsync(g->d + sp.off, x->in[k] + hoff, sizeof(float) * (size_t)sp.n, cudaMemcpyDeviceToDevice, s->stream));
    else MF_LAUNCH(s, k_aos_to_soa, ewBlocks(s, sp.n), 256, x->in[k] + 3 * hoff, g->comp(0) + sp.off, g->comp(1) + sp.off, g->comp(2) + sp.off, sp.n);
  }
  CUDA_CHECK(cudaEventRecord(x->inFree[k], s->stream)); x->inFreeSet[k] = true;
  x->inHead ^= 1; x->inCount--;
  s->rankSync();
  MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_grid_download_async(const mf_grid* gp, void* host) {
  MF_API_BEGIN
  mfAssert(gp && host, "Grid download_async: null argument");
  const Grid* g = G(gp); Solver* s = g->s; s->use();
  GridXfer* x = xferOf(g);
  const int k = x->outNext; x->outNext ^= 1;
  if (!x->out[k]) CUDA_CHECK(cudaMalloc(&x->out[k], x->bytes));
  if (x->outFreeSet[k]) CUDA_CHECK(cudaStreamWaitEvent(s->stream, x->outFree[k], 0));
  const long long base = localSpan(s).off;
  MF_RANKS(s, r) {
    const Span sp = span(s, r);
    if (!sp.n) continue;
    const long long hoff = sp.off - base;
    if (!g->isVec()) CUDA_CHECK(cudaMemcpyAsync(x->out[k] + hoff, g->d + sp.off, sizeof(float) * (size_t)sp.n, cudaMemcpyDeviceToDevice, s->stream));
    else MF_LAUNCH(s, k_soa_to_aos, ewBlocks(s, sp.n), 256, x->out[k] + 3 * hoff, g->comp(0) + sp.off, g->comp(1) + sp.off, g->comp(2) + sp.off, sp.n);
  }
  CUDA_CHECK(cudaEventRecord(x->outReady[k], s->stream));
  CUDA_CHECK(cudaStreamWaitEvent(s->copyOut, x->outReady[k], 0));
  CUDA_CHECK(cudaMemcpyAsync(host, x->out[k], x->bytes, cudaMemcpyDeviceToHost, s->copyOut));
  CUDA_CHECK(cudaEventRecord(x->outFree[k], s->copyOut)); x->outFreeSet[k] = true;
  MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_solver_transfers_sync(mf_solver* sp) {
  MF_API_BEGIN
  mfAssert(sp, "mf_solver_transfers_sync: null solver");
  Solver* s = S(sp); s->use();
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  if (s->copyIn) { CUDA_CHECK(cudaStreamSynchronize(s->copyIn)); CUDA_CHECK(cudaStreamSynchronize(s->copyOut)); }
  MF_API_END
}
int mf_grid_upload(mf_grid* gp, const void* host) { MF_API_BEGIN mfAssert(gp && host, "Grid upload: null argument"); gridUpload(G(gp), host, false); MF_API_END }
int mf_grid_download(const mf_grid* gp, void* host) { MF_API_BEGIN mfAssert(gp && host, "Grid download: null argument"); gridDownload(G(gp), host, false); MF_API_END }
int mf_grid_upload_slab(mf_grid* gp, const void* host) { MF_API_BEGIN mfAssert(gp && host, "Grid upload: null argument"); gridUpload(G(gp), host, true); MF_API_END }
int mf_grid_download_slab(const mf_grid* gp, void* host) { MF_API_BEGIN mfAssert(gp && host, "Grid download: null argument"); gridDownload(G(gp), host, true); MF_API_END }
int mf_grid_set_const_real(mf_grid* gp, float v) {
  MF_API_BEGIN
  REQ_GRID(gp, G(gp)->isReal(), "grid", "Grid<Real>::setConst");
  Grid* g = G(gp); Solver* s = g->s; s->use();
  MF_EW(s, k_fill_f, g->d + sp.off, v, sp.n); s->rankSync(); MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_grid_set_const_int(mf_grid* gp, int v) {
  MF_API_BEGIN
  REQ_GRID(gp, G(gp)->isInt(), "grid", "Grid<int>::setConst");
  Grid* g = G(gp); Solver* s = g->s; s->use(); markFlagsChanged(g);
  MF_EW(s, k_fill_i, g->di() + sp.off, v, sp.n); s->rankSync(); MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_grid_set_const_vec3(mf_grid* gp, float x, float y, float z) {
  MF_API_BEGIN
  REQ_GRID(gp, G(gp)->isVec(), "grid", "Grid<Vec3>::setConst");
  Grid* g = G(gp); Solver* s = g->s; s->use();
  const float f[3] = { x, y, z };
  for (int c = 0; c < 3; c++) MF_EW(s, k_fill_f, g->comp(c) + sp.off, f[c], sp.n);
  s->rankSync(); MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_grid_add_scaled(mf_grid* gp, const mf_grid* ap, float fx, float fy, float fz) {
  MF_API_BEGIN
  mfAssert(gp && ap, "Grid::addScaled: null grid");
  Grid* g = G(gp); const Grid* a = G(ap); Solver* s = g->s; s->use();
  check_same_solver(g, a, "Grid::addScaled");
  mfAssert(!g->isInt() && g->isVec() == a->isVec(), "Grid::addScaled: grid types do not match");
  if (!g->isVec()) { MF_EW(s, k_scaled_add, g->d + sp.off, a->d + sp.off, fx, sp.n); }
  else {
    const float f[3] = { fx, fy, fz };
    for (int c = 0; c < 3; c++) MF_EW(s, k_scaled_add, g->comp(c) + sp.off, a->comp(c) + sp.off, f[c], sp.n);
  }
  s->rankSync(); MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_grid_mult_const(mf_grid* gp, float fx, float fy, float fz) {
  MF_API_BEGIN
  mfAssert(gp, "Grid::multConst: null grid");
  Grid* g = G(gp); Solver* s = g->s; s->use();
  mfAssert(!g->isInt(), "Grid::multConst: int grids not supported");
  if (!g->isVec()) { MF_EW(s, k_mult_const, g->d + sp.off, fx, sp.n); }
  else {
    const float f[3] = { fx, fy, fz };
    for (int c = 0; c < 3; c++) MF_EW(s, k_mult_const, g->comp(c) + sp.off, f[c], sp.n);
  }
  s->rankSync(); MF_LAUNCH_CHECK();
  MF_API_END
}
int mf_grid_max_abs(const mf_grid* gp, float* out) {
  MF_API_BEGIN
  mfAssert(gp && out, "Grid::getMaxAbs: null argument");
  const Grid* g = G(gp); Solver* s = g->s; s->use();
  if (g->isInt()) { int mn, mx; reduce_minmax_int(s, g->di(), s->N, &mn, &mx); *out = (float)(abs(mn) > abs(mx) ? abs(mn) : abs(mx)); }
  else if (!g->isVec()) { float mn, mx; reduce_minmax(s, g->d, s->N, &mn, &mx); *out = fmaxf(fabsf(mn), fabsf(mx)); }
  else {
    // Grid<Vec3>::getMaxAbs = sqrt(max normSquare) (grid.cpp:366-368 via CompMaxVec)
    TmpReal t(s);
    launch_norm_square(s, g, t.p);
    float mn, mx; reduce_minmax(s, t.p, s->N, &mn, &mx); *out = sqrtf(mx);
  }
  MF_API_END
}
int mf_grid_min_max(const mf_grid* gp, float* omin, float* omax) {
  MF_API_BEGIN
  mfAssert(gp && omin && omax, "Grid::getMin/getMax: null argument");
  const Grid* g = G(gp); Solver* s = g->s; s->use();
  mfAssert(!g->isVec(), "Grid::getMin/getMax: Vec3 grids not supported here");
  if (g->isInt()) { int mn, mx; reduce_minmax_int(s, g->di(), s->N, &mn, &mx); *omin = (float)mn; *omax = (float)mx; }
  else reduce_minmax(s, g->d, s->N, omin, omax);
  MF_API_END
}
int mf_grid_sum_sqr(const mf_grid* gp, double* out) {
  MF_API_BEGIN
  REQ_GRID(gp, G(gp)->isReal(), "grid", "GridSumSqr");
  mfAssert(out, "GridSumSqr: null output");
  *out = reduce_sumsqr(G(gp)->s, G(gp)->d, G(gp)->s->N);
  MF_API_END
}
int mf_grid_dot(const mf_grid* ap, const mf_grid* bp, double* out) {
  MF_API_BEGIN
  REQ_GRID(ap, G(ap)->isReal(), "a", "GridDotProduct");
  REQ_GRID(bp, G(bp)->isReal(), "b", "GridDotProduct");
  mfAssert(out, "GridDotProduct: null output");
  check_same_solver(G(ap), G(bp), "GridDotProduct");
  G(ap)->s->use();
  *out = reduce_dot(G(ap)->s, G(ap)->d, G(bp)->d, G(ap)->s->N);
  MF_API_END
}
int mf_grid_max_diff(const mf_grid* ap, const mf_grid* bp, double* out) {
  MF_API_BEGIN
  mfAssert(ap && bp && out, "gridMaxDiff: null argument");
  const Grid* a = G(ap); const Grid* b = G(bp); Solver* s = a->s; s->use();
  check_same_solver(a, b, "gridMaxDiff");
  mfAssert(a->isVec() == b->isVec() && a->isInt() == b->isInt(), "gridMaxDiff: grid types do not match");
  MF_RANKS(s, r) {
    const Span sp = span(s, r);
    if (a->isInt()) { MF_LAUNCH(s, k_maxdiff<int>, redBlocks(s, sp.n), RED_THREADS, a->di() + sp.off, b->di() + sp.off, sp.n, s->Np, 1, redCtx(s, r), s->redOut); }
    else { MF_LAUNCH(s, k_maxdiff<float>, redBlocks(s, sp.n), RED_THREADS, a->d + sp.off, b->d + sp.off, sp.n, s->Np, a->isVec() ? 3 : 1, redCtx(s, r), s->redOut); }
  }
  redFinish<1>(s, RedOps<1>{ { 2 } }, FinStore<1>{ s->redOut });
  MF_LAUNCH_CHECK();
  fetchRed(s, 1);
  *out = s->redHost[0];
  MF_API_END
}

} // extern "C"

namespace mf {
__global__ void k_norm_square(const float* x, const float* y, const float* z, float* o, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) o[i] = x[i] * x[i] + y[i] * y[i] + z[i] * z[i];
}
}
void launch_norm_square(mf::Solver* s, const mf::Grid* g, float* o) {
  using namespace mf;
  MF_EW(s, mf::k_norm_square, g->comp(0) + sp.off, g->comp(1) + sp.off, g->comp(2) + sp.off, o + sp.off, sp.n);
  s->rankSync();
}
