/* mantacuda.h -- C-ABI of libmantacuda.so, the B200 (sm_100a) implementation of Mantaflow's
 * per-step Eulerian grid hot path (advection -> wall BCs / forces -> pressure projection).
 *
 * Every entry point replaces the body of one reference function / PYTHON() plugin; the
 * reference location is cited next to it (paths relative to the reference checkout).
 * INTEGRATION.md shows the plugin-side stub a Mantaflow maintainer would add.
 *
 * Conventions
 *  - plain C types only; grids and solvers are opaque handles owned by the library.
 *  - host buffers passed to upload/download use the REFERENCE layout: x fastest,
 *    idx = i + nx*(j + ny*k); Real = float, int = int32, Vec3 = 3 packed floats (AoS, 12 B)
 *    (source/grid.h:77, source/util/vectorbase.h:202-213).
 *  - every function returns 0 on success, non-zero on error (never throws across the ABI);
 *    mf_last_error() returns the message of the last failure on the calling thread, in the
 *    reference's errMsg style (source/general.h:75-77).
 *  - work is enqueued on the solver's CUDA stream; calls that return values to the host
 *    (download, reductions, solve statistics) synchronise that stream, nothing else does.
 *  - there is NO CPU fallback: without a CUDA device mf_solver_create fails.
 */
#ifndef MANTACUDA_H
#define MANTACUDA_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mf_solver mf_solver;
typedef struct mf_grid   mf_grid;
typedef struct mf_comm   mf_comm;

/* GridBase::GridType bits, source/grid.h:29 */
enum { MF_TYPE_REAL = 1, MF_TYPE_INT = 2, MF_TYPE_VEC3 = 4, MF_TYPE_MAC = 8, MF_TYPE_LEVELSET = 16, MF_TYPE_FLAGS = 32 };
/* FlagGrid::CellType bits, source/grid.h:306-320 */
enum { MF_FLUID = 1, MF_OBSTACLE = 2, MF_EMPTY = 4, MF_INFLOW = 8, MF_OUTFLOW = 16, MF_OPEN = 32, MF_STICK = 64 };
/* enum Preconditioner, source/plugin/pressure.cpp:27 */
enum { MF_PC_NONE = 0, MF_PC_MIC = 1, MF_PC_MG_DYNAMIC = 2, MF_PC_MG_STATIC = 3 };

/* ---- library ---------------------------------------------------------------------- */
int          mf_version(void);
const char*  mf_last_error(void);
int          mf_device_count(void);                      /* number of visible CUDA devices */
const char*  mf_build_info(void);                        /* cf. buildInfoString(), source/general.cpp */

/* ---- FluidSolver: grid size, dt, temp-grid pool (source/fluidsolver.h:27-119, .cpp:25-50,107) */
mf_solver*   mf_solver_create(int nx, int ny, int nz, int dim, int device);
void         mf_solver_destroy(mf_solver* s);
int          mf_solver_set_dt(mf_solver* s, float dt);   /* FluidSolver::mDt ("timestep") */
float        mf_solver_get_dt(const mf_solver* s);
int          mf_solver_sync(mf_solver* s);                /* cudaStreamSynchronize on the solver stream */
void*        mf_solver_stream(mf_solver* s);              /* cudaStream_t, for CUDA-event timing by the caller */
long long    mf_solver_launch_count(const mf_solver* s);  /* kernels launched so far by this library for s */
long long    mf_solver_pool_bytes(const mf_solver* s);    /* device bytes currently held by the temp-grid pool */
/* Per-kernel device timing, the role of the reference's TimingData / MG_TIMINGS (source/timing.cpp:23-112,
 * multigrid.cpp:39-40): between _begin and _end every kernel launch is bracketed by a CUDA-event pair on the solver
 * stream; _end synchronises and writes "kernel launches total_ms\n" lines (sorted by time) into buf. */
int          mf_solver_profile_begin(mf_solver* s);
int          mf_solver_profile_end(mf_solver* s, char* buf, int buflen);
/* CUDA-event stopwatch on the solver stream, slots 0..15 (cf. MuTime, source/general.cpp:51-66) */
int          mf_solver_event_record(mf_solver* s, int slot);
int          mf_solver_event_elapsed_ms(mf_solver* s, int from_slot, int to_slot, float* ms);
/* page-locked host buffers for upload/download (cudaMallocHost); plain malloc'ed memory works too, only slower */
void*        mf_host_alloc(long long bytes);
void         mf_host_free(void* p);

/* ---- multi-GPU: one process per GPU, z-slab partition, one global address space over NVLink ---------------
 * The reference is a shared-memory program whose KERNELs split the z range over TBB/OpenMP threads
 * (source/preprocessor/codegen_kernel.cpp:187-190; source/kernel.h:65).  mf_comm is the B200 counterpart of that thread
 * team: every rank maps every grid into one contiguous virtual range (CUDA VMM + POSIX-fd exchange), hosts one z-slab of
 * it and launches each kernel on its own z range; neighbour planes are peer loads over NVLink, dependent launches are
 * separated by a device-side flag barrier, reductions combine per-rank partials in rank order on every rank.
 * All calls on a distributed solver are COLLECTIVE: every rank makes the same calls in the same order. */
mf_comm*     mf_comm_create(int rank, int nranks, int device, const char* session);  /* session: any name unique to the job */
/* all nranks ranks inside THIS process on one device: every launch group runs rank after rank on one stream, barriers are
 * stream order.  Exercises the partitioned code path (z ranges, cross-rank reductions, solo levels) on a single GPU. */
mf_comm*     mf_comm_create_emulated(int nranks, int device);
void         mf_comm_destroy(mf_comm* c);
/* FluidSolver whose grids are z-slab partitioned over the ranks of c (3-D, nz >= 2*nranks) */
mf_solver*   mf_solver_create_dist(mf_comm* c, int nx, int ny, int nz, int dim);
int          mf_solver_ranks(const mf_solver* s);                              /* 1 for a single-GPU solver */
int          mf_solver_slab(const mf_solver* s, int rank, int* k0, int* k1);   /* planes [k0,k1) of rank (-1: this process) */
long long    mf_solver_barrier_count(const mf_solver* s);                      /* device barriers enqueued so far */
/* host buffers that hold only this rank's slab (planes k0..k1-1, reference layout); mf_grid_upload/_download take the FULL
 * grid on every rank (upload: each rank copies its own slab from it; download: every rank receives the whole grid) */
int          mf_grid_upload_slab(mf_grid* g, const void* host_slab);
int          mf_grid_download_slab(const mf_grid* g, void* host_slab);
/* Streaming transfers for pipelines whose data lives on the host (pinned memory from mf_host_alloc; slab layout as above).
 * Copies run on two copy streams beside the solver stream through double-buffered device staging, so the upload of the next
 * step's inputs and the download of the previous step's results overlap the kernels of the current step:
 *   mf_grid_stage_upload    start the host->device copy of the slab (returns at once; at most 2 staged uploads per grid)
 *   mf_grid_commit_upload   the solver stream waits for the oldest staged copy and moves it into the grid (AoS->SoA for vec grids)
 *   mf_grid_download_async  snapshot the grid on the solver stream, copy it device->host on the copy-out stream
 *   mf_solver_transfers_sync  wait for everything: host buffers are complete / reusable afterwards
 * A host buffer passed to stage_upload / download_async must stay untouched until transfers_sync (or two further calls on the
 * same grid have been followed by a solver sync). */
int          mf_grid_stage_upload(mf_grid* g, const void* pinned_slab);
int          mf_grid_commit_upload(mf_grid* g);
int          mf_grid_download_async(const mf_grid* g, void* pinned_slab);
int          mf_solver_transfers_sync(mf_solver* s);
/* out[0] mismatches seen through peer mappings, out[1] device-barrier latency (us), out[2]/out[3] peer read/write GB/s,
 * out[4] allocation granularity (bytes), out[5] local copy GB/s */
int          mf_comm_selftest(mf_comm* c, long long bytesPerRank, int iters, double* out);

/* ---- Grid<T> / MACGrid / FlagGrid (source/grid.h:99,255,298; grid.cpp:49,95,100,228) - */
mf_grid*     mf_grid_create(mf_solver* s, int type);      /* type = OR of MF_TYPE_*; zero-filled */
void         mf_grid_destroy(mf_grid* g);
int          mf_grid_type(const mf_grid* g);
int          mf_grid_clear(mf_grid* g);                                   /* Grid::clear  grid.cpp:95 */
int          mf_grid_copy(mf_grid* dst, const mf_grid* src);             /* Grid::copyFrom grid.cpp:228 */
int          mf_grid_swap(mf_grid* a, mf_grid* b);                       /* Grid::swap   grid.cpp:100 */
int          mf_grid_get_component(const mf_grid* source, mf_grid* target, int component);   /* getComponent grid.cpp:746 */
int          mf_grid_set_component(const mf_grid* source, mf_grid* target, int component);   /* setComponent grid.cpp:751 */
int          mf_grid_upload(mf_grid* g, const void* host);               /* host -> device, reference layout */
int          mf_grid_download(const mf_grid* g, void* host);             /* device -> host, reference layout */
int          mf_grid_set_const_real(mf_grid* g, float v);                /* Grid::setConst grid.cpp:288 */
int          mf_grid_set_const_int (mf_grid* g, int v);
int          mf_grid_set_const_vec3(mf_grid* g, float x, float y, float z);
int          mf_grid_add_scaled(mf_grid* g, const mf_grid* a, float fx, float fy, float fz); /* Grid::addScaled */
int          mf_grid_mult_const(mf_grid* g, float fx, float fy, float fz);                   /* Grid::multConst */
int          mf_grid_max_abs(const mf_grid* g, float* out);              /* Grid::getMaxAbs grid.cpp:356-368 */
int          mf_grid_min_max(const mf_grid* g, float* out_min, float* out_max); /* getMin/getMax (Real, int) */
int          mf_grid_sum_sqr(const mf_grid* g, double* out);             /* GridSumSqr commonkernels.h:32 */
int          mf_grid_dot(const mf_grid* a, const mf_grid* b, double* out);/* GridDotProduct conjugategrad.cpp:175 */
int          mf_grid_max_diff(const mf_grid* a, const mf_grid* b, double* out); /* gridMaxDiff* grid.cpp:437-467 */

/* ---- allocation-free reduction ---------------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).
 * GridDotProduct's arithmetic (float products summed in double) on caller-owned DEVICE arrays a_dev[n], b_dev[n]; the result is
 * written to *out_dev (device memory) asynchronously on `stream` (a cudaStream_t, NULL = the default stream) of the CURRENT
 * device.  Allocates nothing: the reduction's scratch is a file-scope static __device__ array of MF_NOALLOC_SCRATCH_BYTES that
 * every GPU loading the library holds until exit.  n > MF_NOALLOC_MAX_ELEMS is rejected (status != 0, mf_last_error()).
 * All calls on a device share that scratch: calls that overlap on the GPU race -- serialise them (one stream, or graph edges). */
#define MF_NOALLOC_MAX_ELEMS      33554432LL      /* 4096 blocks x 8192 elements = 2 x 256^3 floats per input */
#define MF_NOALLOC_SCRATCH_BYTES  33024           /* 4096 partial sums (double) + the ticket counter's 256-byte line */
int          mf_dot_product_noalloc(const float* a_dev, const float* b_dev, long long n, double* out_dev, void* stream);

/* FlagGrid setup (source/grid.cpp:798-927, plugin/extforces.cpp:106-131) */
int          mf_flags_init_domain(mf_grid* flags, int boundaryWidth, const char* wall, const char* open,
                                  const char* inflow, const char* outflow, mf_grid* phiWalls /* or NULL */);
int          mf_flags_fill_grid(mf_grid* flags, int type);
int          mf_flags_count_cells(const mf_grid* flags, int flag, int bnd, int* out);
int          mf_flags_set_open_bound(mf_grid* flags, int bWidth, const char* openBound, int type);
/* Box obstacle / source helper: set value (or flag) in cells whose centre lies in [p0,p1]
 * (Shape::applyToGrid with a Box, source/shapes.cpp:71,178) */
int          mf_apply_box_to_real(mf_grid* g, float x0, float y0, float z0, float x1, float y1, float z1, float value);
int          mf_apply_box_to_flags(mf_grid* g, float x0, float y0, float z0, float x1, float y1, float z1, int value);
/* Cylinder::applyToGrid (source/shapes.cpp:71,367): centre c, axis vector z (half-length = |z|), radius r */
int          mf_apply_cylinder_to_real(mf_grid* g, float cx, float cy, float cz, float r, float zx, float zy, float zz, float value);

/* ---- advection (source/plugin/advection.cpp:443) ------------------------------------- */
int          mf_advect_semi_lagrange(const mf_grid* flags, const mf_grid* vel, mf_grid* grid,
                                     int order, float strength, int orderSpace, int clampMode, int orderTrace);

/* ---- boundary conditions and forces (source/plugin/extforces.cpp) -------------------- */
int          mf_set_wall_bcs(const mf_grid* flags, mf_grid* vel, const mf_grid* obvel, const mf_grid* fractions,
                             const mf_grid* phiObs, int boundaryWidth);                     /* :327 */
int          mf_add_buoyancy(const mf_grid* flags, const mf_grid* density, mf_grid* vel,
                             float gx, float gy, float gz, float coefficient, int scale);   /* :84  */
int          mf_add_gravity(const mf_grid* flags, mf_grid* vel, float gx, float gy, float gz,
                            const mf_grid* exclude, int scale);                             /* :62  */
int          mf_set_inflow_bcs(mf_grid* vel, const char* dir, float vx, float vy, float vz);/* :171 */
int          mf_reset_outflow(mf_grid* flags, mf_grid* phi, mf_grid* real);                 /* :134 (grid part) */

/* ---- pressure projection (source/plugin/pressure.cpp) -------------------------------- */
typedef struct mf_pressure_params {          /* keyword arguments of solvePressure, pressure.cpp:482-497 */
  float  cgAccuracy;            /* 1e-3 */
  const mf_grid* phi;           /* 0 */
  const mf_grid* perCellCorr;   /* 0 */
  const mf_grid* fractions;     /* 0 */
  const mf_grid* obvel;         /* 0 */
  float  gfClamp;               /* 1e-4 */
  float  cgMaxIterFac;          /* 1.5 */
  int    precondition;          /* true (deprecated switch) */
  int    preconditioner;        /* MF_PC_MIC */
  int    enforceCompatibility;  /* false */
  int    useL2Norm;             /* false */
  int    zeroPressureFixing;    /* false */
  const mf_grid* curv;          /* NULL */
  float  surfTens;              /* 0 */
  mf_grid* retRhs;              /* NULL (solvePressure only) */
} mf_pressure_params;
void         mf_pressure_params_default(mf_pressure_params* p);

typedef struct mf_solve_info {               /* what the reference prints at debug level 2, pressure.cpp:442 */
  int    iterations;
  float  resNorm;
  int    maxIter;
  int    fixedCell;             /* linear index pinned by zeroPressureFixing, or -1 */
  int    mgLevels;              /* 0 if no multigrid */
  int    mgRebuilt;             /* 1 if the hierarchy was (re)built in this call */
} mf_solve_info;

int          mf_compute_pressure_rhs(mf_grid* rhs, const mf_grid* vel, const mf_grid* pressure, const mf_grid* flags,
                                     const mf_pressure_params* p);                           /* :277 */
int          mf_solve_pressure_system(mf_grid* rhs, mf_grid* vel, mf_grid* pressure, const mf_grid* flags,
                                      const mf_pressure_params* p, mf_solve_info* info);     /* :312 */
int          mf_correct_velocity(mf_grid* vel, mf_grid* pressure, const mf_grid* flags,
                                 const mf_pressure_params* p);                               /* :457 */
int          mf_solve_pressure(mf_grid* vel, mf_grid* pressure, const mf_grid* flags,
                               const mf_pressure_params* p, mf_solve_info* info);            /* :482 */
int          mf_release_mg(mf_solver* s);                                                    /* :252 */

/* ---- scene-setup plugins that run inside the step loop -- SURVEY 8(f) rank 1 ---------------------------------------------- */
enum { MF_SHAPE_BOX = 0, MF_SHAPE_SPHERE = 1, MF_SHAPE_CYLINDER = 2 };
/* Box: a = p0, b = p1.  Sphere: a = center, b = scale, r = radius.  Cylinder: a = center, b = z (axis, its length is the half
 * height, normalised by the library as Cylinder::Cylinder does), r = radius.          source/shapes.h:93,120,139 */
typedef struct mf_shape { int type; float a[3]; float b[3]; float r; } mf_shape;
/* Shape::applyToGrid(grid, value, respectFlags=NULL), shapes.cpp:72: int grids take ivalue, Real grids vx, Vec3 / MAC grids
 * (vx, vy, vz) */
int          mf_shape_apply(const mf_shape* shape, mf_grid* grid, float vx, float vy, float vz, int ivalue, const mf_grid* respectFlags);
/* Shape::computeLevelset() / generateLevelset(phi), shapes.cpp:29 (BoxSDF :178, SphereSDF :309, CylinderSDF :367) */
int          mf_shape_levelset(const mf_shape* shape, mf_grid* phi);
int          mf_flags_update_from_levelset(mf_grid* flags, const mf_grid* levelset);     /* grid.cpp:910 */
int          mf_levelset_join(mf_grid* a, const mf_grid* b);                             /* levelset.cpp:111 */
/* updateFractions(flags, phiObs, fractions, boundaryWidth=0, fracThreshold=0.01), plugin/initplugins.cpp:437 */
int          mf_update_fractions(const mf_grid* flags, const mf_grid* phiObs, mf_grid* fractions, int boundaryWidth, float fracThreshold);
/* setObstacleFlags(flags, phiObs, fractions=NULL, phiOut=NULL, phiIn=NULL, boundaryWidth=1), plugin/initplugins.cpp:473 */
int          mf_set_obstacle_flags(mf_grid* flags, const mf_grid* phiObs, const mf_grid* fractions, const mf_grid* phiOut,
                                   const mf_grid* phiIn, int boundaryWidth);

/* ---- simple extrapolation (source/fastmarch.cpp) -- SURVEY 8(f) rank 2 ------------------------------------------------ */
/* extrapolateMACSimple(flags, vel, distance=4, phiObs=NULL, intoObs=false), fastmarch.cpp:337 */
int          mf_extrapolate_mac_simple(const mf_grid* flags, mf_grid* vel, int distance, const mf_grid* phiObs, int intoObs);
/* extrapolateMACFromWeight(vel, weight, distance=2), fastmarch.cpp:410 (the weight grid is overwritten) */
int          mf_extrapolate_mac_from_weight(mf_grid* vel, mf_grid* weight, int distance);
/* extrapolateLsSimple(phi, distance=4, inside=false, include_walls=false), fastmarch.cpp:471 */
int          mf_extrapolate_ls_simple(mf_grid* phi, int distance, int inside, int include_walls);
/* extrapolateVec3Simple(vel, phi, distance=4, inside=false), fastmarch.cpp:525 */
int          mf_extrapolate_vec3_simple(mf_grid* vel, const mf_grid* phi, int distance, int inside);

/* pieces of the solver exposed for kernel-level parity tests */
int          mf_make_laplace_matrix(const mf_grid* flags, mf_grid* A0, mf_grid* Ai, mf_grid* Aj, mf_grid* Ak,
                                    const mf_grid* fractions);          /* conjugategrad.h:154 */
int          mf_apply_matrix(const mf_grid* flags, mf_grid* dst, const mf_grid* src,
                             const mf_grid* A0, const mf_grid* Ai, const mf_grid* Aj, const mf_grid* Ak); /* conjugategrad.h:118/136 */
/* dst = A(flags) * src through the CG's 16-bit row codes (MakeLaplaceMatrix without fractions + ApplyMatrix fused; needs
   nx % 4 == 0): must equal mf_make_laplace_matrix + mf_apply_matrix bit for bit. */
int          mf_apply_laplace_rowcode(const mf_grid* flags, mf_grid* dst, const mf_grid* src);
/* ---- GridCg as an object: the interface the reference's C++-side clients drive directly (GridCgInterface, conjugategrad.h:27-58;
   callers: pressure.cpp:372-452, fluidguiding.cpp:276-334, waves.cpp:129-133, vortexplugins.cpp:279).
   mf_cg_create = GridCg<ApplyMatrix>(dst, rhs, residual, search, flags, tmp, A0, Ai, Aj, Ak) (conjugategrad.cpp:201; ApplyMatrix2D on a
   2-D solver): all grids stay owned by the caller and must outlive the object; accuracy starts at VECTOR_EPSILON (1e-6), the norm
   is the L2 one (GridCgInterface() :31) until mf_cg_set_use_l2_norm(cg, 0).  iterate / solve follow conjugategrad.cpp:238-308:
   the first call initialises (dst = 0, residual = rhs, preconditioner set-up, sigma); *notConverged = iterate()'s return value.
   method: 0 PC_None, 1 PC_ICP, 2 PC_mICP (both run the red-black modified IC(0) of this library, 3-D only as in the reference:
   silently disabled in 2-D, conjugategrad.cpp:315-321), 3 PC_MGP (the solver's cached GridMg -- gMapMG, pressure.cpp:250 -- is
   created / set up from A0..Ak on first use; rebuild != 0 drops a cached hierarchy first, like passing a fresh GridMg). */
typedef struct mf_cg mf_cg;
mf_cg*       mf_cg_create(mf_grid* dst, mf_grid* rhs, mf_grid* residual, mf_grid* search, const mf_grid* flags, mf_grid* tmp,
                          mf_grid* A0, mf_grid* Ai, mf_grid* Aj, mf_grid* Ak);
void         mf_cg_destroy(mf_cg* cg);
int          mf_cg_set_ic_preconditioner(mf_cg* cg, int method);                 /* setICPreconditioner, conjugategrad.cpp:311 */
int          mf_cg_set_mg_preconditioner(mf_cg* cg, int method, int rebuild);    /* setMGPreconditioner, conjugategrad.cpp:329 */
int          mf_cg_set_accuracy(mf_cg* cg, float accuracy);                      /* setAccuracy */
int          mf_cg_set_use_l2_norm(mf_cg* cg, int useL2);                        /* setUseL2Norm */
int          mf_cg_force_reinit(mf_cg* cg);                                      /* forceReinit */
int          mf_cg_iterate(mf_cg* cg, int* notConverged);                        /* iterate, conjugategrad.cpp:238 */
int          mf_cg_solve(mf_cg* cg, int maxIter);                                /* solve, conjugategrad.cpp:302 */
int          mf_cg_get(mf_cg* cg, float* sigma, float* iterations, float* resNorm, float* accuracy);   /* getSigma / getIterations / getResNorm / getAccuracy */
/* cgSolveDiffusion(flags, grid, alpha, cgMaxIterFac, cgAccuracy), conjugategrad.cpp:350 */
int          mf_cg_solve_diffusion(const mf_grid* flags, mf_grid* grid, float alpha, float cgMaxIterFac, float cgAccuracy);

/* ---- wavelet noise and noise-driven sources (source/noisefield.h:26-97, source/plugin/initplugins.cpp:26-50) ---- */
typedef struct mf_noise mf_noise;
/* the python-visible members of WaveletNoiseField (noisefield.h:49-60); passed with every call, like attribute reads */
typedef struct mf_noise_params {
  float posOffset[3]; float posScale[3];
  float valOffset; float valScale;
  int clamp; float clampNeg; float clampPos;
  float timeAnim;
} mf_noise_params;
/* WaveletNoiseField(parent, fixedSeed=-1, loadFromFile=false), noisefield.cpp:65-84: seed offset from the seeded stream, the
 * 3 x 128^3 tile generated once per process (or read from / written to ./waveletNoiseTile.bin when loadFromFile) */
mf_noise*    mf_noise_create(mf_solver* s, int fixedSeed, int loadFromFile);
void         mf_noise_destroy(mf_noise* n);
/* host-only: WaveletNoiseField::generateTile (noisefield.cpp:97-181) into out[3 * 128^3]; no device needed */
int          mf_noise_generate_tile(float* out);
int          mf_noise_seed_offset(const mf_noise* n, float* out3);
/* WaveletNoiseField::evaluate(pos) at `count` host positions (xyz interleaved) -> host values; solverTime = FluidSolver::getTime() */
int          mf_noise_evaluate(const mf_noise* n, const mf_noise_params* p, float solverTime, const float* pos, int count, float* out);
/* densityInflow(flags, density, noise, shape, scale=1, sigma=0), initplugins.cpp:39 */
int          mf_density_inflow(const mf_grid* flags, mf_grid* density, const mf_noise* n, const mf_noise_params* p, const mf_shape* shape,
                               float scale, float sigma, float solverTime);
/* addNoise(flags, density, noise, sdf=NULL, scale=1), initplugins.cpp:49 */
int          mf_add_noise(const mf_grid* flags, mf_grid* density, const mf_noise* n, const mf_noise_params* p, const mf_grid* sdf, float scale,
                          float solverTime);

/* ---- FLIP particle system (source/particle.h:32-180,198; source/plugin/flip.cpp) ---- */
typedef struct mf_parts mf_parts;
enum { MF_INT_EULER = 0, MF_INT_RK2 = 1, MF_INT_RK4 = 2 };            /* IntegrationMode, util/integrator.h:23 */
mf_parts*    mf_parts_create(mf_solver* s);                             /* BasicParticleSystem(parent), particle.h:200 */
void         mf_parts_destroy(mf_parts* p);
long long    mf_parts_size(const mf_parts* p);
/* ParticleSystem::create(PdataReal | PdataVec3 | PdataInt): ncomp 1, 3 or -1; *index identifies the data in later calls */
int          mf_parts_create_pdata(mf_parts* p, int ncomp, int* index);
/* host <-> device: positions n x 3 interleaved, flags n ints (ParticleBase::ParticleStatus), particle data n x ncomp */
int          mf_parts_upload(mf_parts* p, const float* pos, const int* flag, long long n);
int          mf_parts_download(const mf_parts* p, float* pos, int* flag);
int          mf_pdata_upload(mf_parts* p, int pdata, const void* in);
int          mf_pdata_download(const mf_parts* p, int pdata, void* out);
/* sampleFlagsWithParticles(flags, parts, discretization, randomness), flip.cpp:32 */
int          mf_sample_flags_with_particles(const mf_grid* flags, mf_parts* parts, int discretization, float randomness);
/* parts.advectInGrid(flags, vel, integrationMode, deleteInObstacle, stopInObstacle, skipNew), particle.h:527 */
int          mf_parts_advect_in_grid(mf_parts* parts, const mf_grid* flags, const mf_grid* vel, int integrationMode, int deleteInObstacle,
                                     int stopInObstacle, int skipNew);
/* mapPartsToMAC(flags, vel, velOld, parts, partVel, weight=NULL), flip.cpp:637 */
int          mf_map_parts_to_mac(const mf_grid* flags, mf_grid* vel, mf_grid* velOld, const mf_parts* parts, int partVel, mf_grid* weight);
/* markFluidCells(parts, flags, phiObs=NULL), flip.cpp:166 */
int          mf_mark_fluid_cells(const mf_parts* parts, mf_grid* flags, const mf_grid* phiObs);
/* mapMACToParts(flags, vel, parts, partVel), flip.cpp:717; flipVelocityUpdate(flags, vel, velOld, parts, partVel, flipRatio), flip.cpp:738 */
int          mf_map_mac_to_parts(const mf_grid* flags, const mf_grid* vel, const mf_parts* parts, int partVel);
int          mf_flip_velocity_update(const mf_grid* flags, const mf_grid* vel, const mf_grid* velOld, const mf_parts* parts, int partVel, float flipRatio);

/* ---- GridMg inspection hooks for parity tests (source/multigrid.h:31-137 keeps these private) ---- */
/* levels / host-fallback count of the solver's cached hierarchy (the gMapMG entry, pressure.cpp:250) */
int          mf_mg_info(mf_solver* s, int* levels, int* usedHostFallback);
/* size3[0..2] = vertex counts of level l; returns non-zero if there is no such level */
int          mf_mg_level_size(mf_solver* s, int l, int* size3);
/* *bytesPerEntry = storage width of level l's operator on the device: 2 (level 0: 16-bit row code; level 1: binary16 when every
 * entry converts exactly) or 4 (float) */
int          mf_mg_level_format(mf_solver* s, int l, int* bytesPerEntry);
/* type: n bytes of GridMg::VertexType; A: n*stencil floats in the reference's interleaved order mA[l][v*stencil+e] */
int          mf_mg_download_level(mf_solver* s, int l, void* type, void* A);
/* GridMg(size); setA(A0..Ak); setCoarsestLevelAccuracy(acc); setSmoothing(1,1); setRhs(rhs); doVCycle(dst)
 * -- the hierarchy stays cached on the solver like PcMGStatic's; A0 == NULL re-applies the cached hierarchy (setRhs + doVCycle only) */
int          mf_mg_vcycle(mf_solver* s, const mf_grid* A0, const mf_grid* Ai, const mf_grid* Aj, const mf_grid* Ak,
                          const mf_grid* rhs, mf_grid* dst, float coarsestAccuracy);

#ifdef __cplusplus
}
#endif
#endif
