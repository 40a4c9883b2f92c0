// common.cuh -- internal data model of libmantacuda (not part of the ABI).
//
// Device data layout (DESIGN.md "Data layout in HBM"):
//   Real / int grids : dense x-fastest array, idx = i + nx*(j + ny*k)  (== reference, grid.h:77)
//   Vec3 / MAC grids : SoA, three component planes of Np floats each (Np = N rounded up to 32
//                      elements so every plane is 128 B aligned).  The reference's 12-byte AoS
//                      (vectorbase.h:202-213) only exists at the upload/download boundary.
// All kernels are compiled with -fmad=false so float arithmetic rounds exactly like the
// reference's x86-64 (non-FMA) build.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <cstdint>
#include <string>
#include <vector>
#include <map>
#include <stdexcept>
#include <sstream>
#include "../../include/mantacuda.h"

namespace mf {

struct Error : public std::runtime_error { Error(const std::string& s) : std::runtime_error(s) {} };

void set_last_error(const std::string& s);

#define MF_STR2(x) #x
#define MF_STR(x) MF_STR2(x)
#define mfErr(msg) do { std::ostringstream _s; _s << msg << "\nError raised in " << __FILE__ << ":" << __LINE__; throw mf::Error(_s.str()); } while (0)
#define mfAssert(cond, msg) do { if (!(cond)) mfErr(msg); } while (0)
#define CUDA_CHECK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) mfErr("CUDA error: " << cudaGetErrorString(_e) << " in " #x); } while (0)

// boundary wrapper: run body, translate exceptions into status codes
#define MF_API_BEGIN try {
#define MF_API_END   return 0; } catch (const std::exception& e) { mf::set_last_error(e.what()); return 1; } catch (...) { mf::set_last_error("unknown error"); return 2; }

struct Solver;
struct MgHierarchy;
struct MicPlan;
struct Comm;
struct CommShared;

struct GridXfer;      // staging of the streaming transfers (solver.cu), lazily created
struct Grid {
  Solver* s;
  int type;          // MF_TYPE_* bits
  float* d;          // device storage (int grids reinterpret)
  bool pooled;       // storage came from the solver pool
  mutable int outflowState;  // flag grids: -1 unknown, 0 no TypeOutflow cell, 1 some (cache for applyOutflowBC)
  mutable unsigned long long version;   // changes whenever an entry point may have written the grid's cells as flags (markFlagsChanged)
  mutable GridXfer* xfer;    // owned by the grid created through mf_grid_create (temporary Grid copies never touch it)
  bool isVec() const { return (type & MF_TYPE_VEC3) != 0; }
  bool isMAC() const { return (type & MF_TYPE_MAC) != 0; }
  bool isReal() const { return (type & MF_TYPE_REAL) != 0; }
  bool isInt() const { return (type & MF_TYPE_INT) != 0; }
  int* di() const { return reinterpret_cast<int*>(d); }
  inline float* comp(int c) const;
};

struct Solver {
  int nx, ny, nz, dim;
  bool is3D;
  long long N;        // cells
  long long Np;       // padded plane stride for SoA vec grids
  long long strideZ;  // nx*ny in 3-D, 0 in 2-D  (grid.cpp:56)
  float dt, dx;
  int device;
  cudaStream_t stream;
  cudaStream_t copyIn, copyOut;   // streaming transfers (mf_grid_stage_upload / _download_async), lazily created
  long long launches;
  // temp-grid pool with the reference's stack discipline (fluidsolver.cpp:33-50)
  std::vector<float*> poolReal, poolVec;
  long long poolBytes;
  // reduction scratch (device) + pinned host result
  double* redScratch;   // per-block partials
  unsigned int* redCounter;
  double* redOut;       // device results (16 doubles)
  double* redHost;      // pinned
  int* devFlagsScratch; // small int scratch (16 ints), job-wide when partitioned over GPUs: any rank's kernels raise flags here
  int* localInts;       // small int scratch (16 ints), private to this process
  volatile unsigned* seqHost; unsigned seqTicket;   // pinned ticket readSmall spins on (solver.cu)
  int* hostInts;        // pinned
  float* aosStage;      // staging buffer for AoS<->SoA (3*N floats), lazily allocated
  MgHierarchy* mg;      // static multigrid cache (gMapMG, pressure.cpp:250)
  MicPlan* mic;         // tile order / flags of the MIC(0) wavefront sweeps (mic.cuh), lazily created
  float* noiseTile;         // 3 x 128^3 wavelet-noise tile (noise.cu), uploaded on first use
  unsigned short* cgCode;   // 16-bit Laplace row codes for the CG's ApplyMatrix (pressure_kernels.cuh CGC_*), lazily allocated
  // the row codes (and the pinned cell of zeroPressureFixing) depend on the flags only: kept while the same flag storage holds
  // the same version (single-GPU solvers)
  const void* codeFlags; unsigned long long codeVersion; int codeFixing; int codeFixedCell;
  int numSMs;
  // ---- z-slab partition over GPUs (comm.h).  Single GPU: nranks = 1, comm = nullptr.
  Comm* comm;           // not owned
  int nranks;           // ranks of the job
  int rankBegin, rankEnd;   // ranks this process executes: [rank, rank+1) per process; [0, nranks) when emulated on one GPU
  int zsplit[17];       // rank r owns planes [zsplit[r], zsplit[r+1])
  int redParity;        // alternates the cross-rank reduction slot set
  bool multi() const { return nranks > 1; }
  // global (all ranks map it) when partitioned over GPUs, else cudaMalloc; zero-filled.  elemSplit[r] = first element (of
  // elemSize bytes) of rank r's share (nranks+1 entries); nullptr = the z-slabs of a cell array with elemSize-byte cells
  void* galloc(size_t bytesPerComp, int ncomp = 1, size_t compStride = 0, const long long* elemSplit = nullptr, int elemSize = 4);
  void  gfree(void* p);
  void  rankSync();      // device barrier across ranks on the solver stream (no-op on a single GPU / emulated)
  void  clearShared(void* p, size_t bytes);   // zero a small job-wide scratch array (one rank clears, then barrier)
  // small control transfers on the solver stream WITHOUT the copy engines (a 40-byte readback must not queue behind a bulk
  // host copy of the streaming transfers): a one-warp kernel stores into mapped pinned memory / carries the payload as argument
  void  readSmall(void* pinnedDst, const void* dev, int bytes);     // then synchronises the stream; bytes % 4 == 0, <= 256
  unsigned readSmallLaunch(void* pinnedDst, const void* dev, int bytes);   // the copy alone; the data is valid after readSmallWait(ticket)
  void  readSmallWait(unsigned ticket);
  void  writeSmall(void* dev, const void* hostSrc, int bytes);      // asynchronous; bytes % 4 == 0, <= 64
  // per-kernel CUDA-event timing (mf_solver_profile_*): one event pair per launch while enabled
  bool profiling;
  struct ProfRec { const char* name; cudaEvent_t a, b; int tag; };
  bool profLevels;      // MF_PROFILE_LEVELS set: split the profile of multigrid kernels by level
  int profTag;          // appended to the kernel name in the profile when >= 0 (multigrid level)
  std::vector<ProfRec> profRecs;
  std::vector<cudaEvent_t> evPool;
  cudaEvent_t profEvent();
  cudaEvent_t userEvents[16];

  float* getReal();           // un-cleared scratch of N floats
  void   putReal(float* p);
  float* getVec();            // un-cleared scratch of 3*Np floats
  void   putVec(float* p);
  void   use() const;         // cudaSetDevice
};

inline float* Grid::comp(int c) const { return d + (long long)c * s->Np; }
// every entry point that writes a flag (int) grid calls this: drops the caches derived from the flags (outflow scan of
// applyOutflowBC, the CG's Laplace row codes and flag statistics of solvePressureSystem)
unsigned long long nextGridVersion();
inline void markFlagsChanged(const Grid* g) { g->outflowState = -1; g->version = nextGridVersion(); }

// RAII temp grids
struct TmpReal { Solver* s; float* p; TmpReal(Solver* s_) : s(s_), p(s_->getReal()) {} ~TmpReal() { s->putReal(p); } operator float*() const { return p; } };
struct TmpVec  { Solver* s; float* p; TmpVec(Solver* s_) : s(s_), p(s_->getVec()) {} ~TmpVec() { s->putVec(p); }  float* comp(int c) const { return p + (long long)c * s->Np; } };

// grid geometry passed by value to kernels
struct Geo {
  int nx, ny, nz;
  int is3D;
  long long Z;   // strideZ (0 in 2-D)
  long long N;
  int k0, k1;    // z range the launch works on ([0,nz) on a single GPU): kernels index memory globally, iterate locally
};
inline Geo geo(const Solver* s) { Geo g; g.nx = s->nx; g.ny = s->ny; g.nz = s->nz; g.is3D = s->is3D; g.Z = s->strideZ; g.N = s->N; g.k0 = 0; g.k1 = s->nz; return g; }
inline Geo geo(const Solver* s, int rank) { Geo g = geo(s); g.k0 = s->zsplit[rank]; g.k1 = s->zsplit[rank + 1]; return g; }
// cell range [off, off+n) of rank r's slab
struct Span { long long off, n; };
inline Span span(const Solver* s, int rank) { Span sp; const long long xy = (long long)s->nx * s->ny; sp.off = xy * s->zsplit[rank]; sp.n = xy * (s->zsplit[rank + 1] - s->zsplit[rank]); return sp; }

// three component pointers of a SoA vec grid
struct V3 { float* x; float* y; float* z; };
struct CV3 { const float* x; const float* y; const float* z; };
inline V3 v3(const Grid* g) { V3 v; v.x = g->comp(0); v.y = g->comp(1); v.z = g->comp(2); return v; }
inline CV3 cv3(const Grid* g) { CV3 v; v.x = g->comp(0); v.y = g->comp(1); v.z = g->comp(2); return v; }
inline CV3 cv3(const TmpVec& t) { CV3 v; v.x = t.comp(0); v.y = t.comp(1); v.z = t.comp(2); return v; }
inline V3 v3(const TmpVec& t) { V3 v; v.x = t.comp(0); v.y = t.comp(1); v.z = t.comp(2); return v; }

inline const Grid* G(const mf_grid* g) { return reinterpret_cast<const Grid*>(g); }
inline Grid* G(mf_grid* g) { return reinterpret_cast<Grid*>(g); }
inline Solver* S(mf_solver* s) { return reinterpret_cast<Solver*>(s); }
inline const Solver* S(const mf_solver* s) { return reinterpret_cast<const Solver*>(s); }

void check_same_solver(const Grid* a, const Grid* b, const char* what);
void release_mg(Solver* s);   // drop the cached multigrid hierarchy (pressure.cu)
void release_mic(Solver* s);  // drop the MIC(0) sweep plan (pressure.cu)
#define REQ_GRID(g, cond, name, fn) mfAssert((g) != nullptr && (cond), fn ": argument '" name "' has the wrong grid type or is missing")

// launch helpers -----------------------------------------------------------------------
inline unsigned int cdiv(long long a, long long b) { return (unsigned int)((a + b - 1) / b); }
struct ProfScope {   // brackets one launch with events on the solver stream when profiling is on
  Solver* s; cudaEvent_t b;
  ProfScope(Solver* s_, const char* name) : s(s_), b(nullptr) {
    if (!s->profiling) return;
    Solver::ProfRec r; r.name = name; r.a = s->profEvent(); r.b = s->profEvent(); b = r.b; r.tag = s->profTag;
    cudaEventRecord(r.a, s->stream); s->profRecs.push_back(r);
  }
  ~ProfScope() { if (b) cudaEventRecord(b, s->stream); }
};
// MF_DEBUG_SYNC=1 (development): synchronise after every launch and name the kernel an error comes from
inline void dbgSync(Solver* s, const char* name) {
  static const bool on = getenv("MF_DEBUG_SYNC") != nullptr;
  if (!on) return;
  cudaError_t e = cudaStreamSynchronize(s->stream);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) fprintf(stderr, "MF_DEBUG_SYNC: %s after %s\n", cudaGetErrorString(e), name);
}
#define MF_LAUNCH(s, kern, grid, block, ...) do { { mf::ProfScope _ps((s), #kern); kern<<<(grid), (block), 0, (s)->stream>>>(__VA_ARGS__); (s)->launches++; } mf::dbgSync((s), #kern); } while (0)
// Programmatic dependent launch (sm_90+): a kernel launched with MF_LAUNCH_PDL may be scheduled while its predecessor on the
// stream is still draining; it must call pdlWait() before it touches global memory (waits for the predecessor's completion and
// memory flush) and calls pdlTrigger() so that ITS successor can be scheduled early.  Hides the ~4 us launch gap between the
// hundreds of small dependent launches of a V-cycle.  Kernels launched with plain MF_LAUNCH keep full stream ordering.
template <class... KArgs, class... Args>
inline void launchPdlS(cudaStream_t st, size_t smem, void (*kern)(KArgs...), dim3 grid, dim3 block, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
template <class... KArgs, class... Args>
inline void launchPdl(cudaStream_t st, void (*kern)(KArgs...), dim3 grid, dim3 block, Args... args) { launchPdlS(st, 0, kern, grid, block, args...); }
// with dynamic shared memory (profiling mode brackets the launch with events like MF_LAUNCH)
#define MF_LAUNCH_PDL_SMEM(s, smem, kern, grid, block, ...) do { { mf::ProfScope _ps((s), #kern); \
  mf::launchPdlS((s)->stream, smem, kern, dim3(grid), dim3(block), __VA_ARGS__); (s)->launches++; } mf::dbgSync((s), #kern); } while (0)
#define MF_LAUNCH_PDL(s, kern, grid, block, ...) do { if ((s)->profiling) { MF_LAUNCH(s, kern, grid, block, __VA_ARGS__); } \
  else { mf::launchPdl((s)->stream, kern, dim3(grid), dim3(block), __VA_ARGS__); (s)->launches++; mf::dbgSync((s), #kern); } } while (0)
#define MF_LAUNCH_CHECK() CUDA_CHECK(cudaGetLastError())
// SPMD launch helpers: MF_RANKS iterates over the ranks this process executes; every group of per-rank launches is followed
// by rankSync (dependent launches on other GPUs read/write the same global arrays).
#define MF_RANKS(s, r) for (int r = (s)->rankBegin; r < (s)->rankEnd; ++r)
inline dim3 ijkBlockOf(const Solver* s) { return s->is3D ? dim3(32, 4, 2) : dim3(32, 8, 1); }
inline dim3 ijkGridOf(dim3 b, const Geo& g) { return dim3(cdiv(g.nx, b.x), cdiv(g.ny, b.y), cdiv(g.k1 - g.k0, b.z)); }
// ijk kernel (first parameter Geo) over every local rank's z range, then the cross-rank barrier
#define MF_LAUNCH_IJK(s, kern, ...) do { const dim3 _b = mf::ijkBlockOf(s); MF_RANKS(s, _r) { const mf::Geo _g = mf::geo((s), _r); \
  if (_g.k1 > _g.k0) { MF_LAUNCH(s, kern, mf::ijkGridOf(_b, _g), _b, _g, __VA_ARGS__); } } (s)->rankSync(); } while (0)

// the same with programmatic dependent launch: for kernels that begin with MF_PDL_PROLOGUE (MF_NO_STEP_PDL=1: plain launches)
inline bool stepPdl() { static const bool on = getenv("MF_NO_STEP_PDL") == nullptr; return on; }
#define MF_LAUNCH_STEP(s, kern, grid, block, ...) do { if (mf::stepPdl()) { MF_LAUNCH_PDL(s, kern, grid, block, __VA_ARGS__); } else { MF_LAUNCH(s, kern, grid, block, __VA_ARGS__); } } while (0)
#define MF_LAUNCH_IJK_PDL(s, kern, ...) do { const dim3 _b = mf::ijkBlockOf(s); MF_RANKS(s, _r) { const mf::Geo _g = mf::geo((s), _r); \
  if (_g.k1 > _g.k0) { MF_LAUNCH_STEP(s, kern, mf::ijkGridOf(_b, _g), _b, _g, __VA_ARGS__); } } (s)->rankSync(); } while (0)

// reductions (reduce.cu)
// Deterministic two-stage reductions: per-block partial in fixed order, last block sums partials in block order.
double reduce_dot(Solver* s, const float* a, const float* b, long long n);     // sum of float(a*b) in double (GridDotProduct)
double reduce_sumsqr(Solver* s, const float* a, long long n);                  // sum of double(a)^2 (GridSumSqr)
void   reduce_minmax(Solver* s, const float* a, long long n, float* mn, float* mx);
void   reduce_minmax_int(Solver* s, const int* a, long long n, int* mn, int* mx);

// device-side helpers ------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ bool fIsFluid(int f) { return (f & MF_FLUID) != 0; }
__device__ __forceinline__ bool fIsObstacle(int f) { return (f & MF_OBSTACLE) != 0; }
__device__ __forceinline__ bool fIsEmpty(int f) { return (f & MF_EMPTY) != 0; }
__device__ __forceinline__ bool fIsOutflow(int f) { return (f & MF_OUTFLOW) != 0; }
__device__ __forceinline__ bool fIsStick(int f) { return (f & MF_STICK) != 0; }

__device__ __forceinline__ void pdlWait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdlTrigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#define MF_PDL_PROLOGUE pdlTrigger(); pdlWait();
__device__ __forceinline__ float warpSum(float v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }
__device__ __forceinline__ double warpSumD(double v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }
#endif

} // namespace mf
