// pressure.cu -- pressure projection orchestration (reference: source/plugin/pressure.cpp:277-523,
// source/conjugategrad.cpp:201-335,350-423).
#include "common.cuh"
#include "reduce.cuh"
#include "pressure_kernels.cuh"
#include "cg.cuh"
#include "mic.cuh"
#include "multigrid.h"
#include <cmath>
#include <cstdlib>

namespace mf {

// slab-wise whole-array operations (each rank touches its own z-slab of the global array)
static void slabZero(Solver* s, float* p) {
  MF_RANKS(s, r) { const Span sp = span(s, r); if (sp.n) CUDA_CHECK(cudaMemsetAsync(p + sp.off, 0, sizeof(float) * sp.n, s->stream)); }
  s->rankSync();
}
static void slabCopy(Solver* s, float* dst, const float* src) {
  MF_RANKS(s, r) { const Span sp = span(s, r); if (sp.n) CUDA_CHECK(cudaMemcpyAsync(dst + sp.off, src + sp.off, sizeof(float) * sp.n, cudaMemcpyDeviceToDevice, s->stream)); }
  s->rankSync();
}
static inline int ewBlocks(const Solver* s, long long n) {
  long long b = (n + 255) / 256, cap = (long long)s->numSMs * 16;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

static void checkParams(const Solver* s, const mf_pressure_params* p, const char* fn) {
  mfAssert(p, fn << ": null params");
  auto chkReal = [&](const mf_grid* g, const char* name) { if (g) { mfAssert(G(g)->isReal(), fn << ": argument '" << name << "' must be a RealGrid"); mfAssert(G(g)->s == s, fn << ": grids belong to different solvers"); } };
  auto chkMac = [&](const mf_grid* g, const char* name) { if (g) { mfAssert(G(g)->isMAC(), fn << ": argument '" << name << "' must be a MACGrid"); mfAssert(G(g)->s == s, fn << ": grids belong to different solvers"); } };
  chkReal(p->phi, "phi"); chkReal(p->perCellCorr, "perCellCorr"); chkReal(p->curv, "curv"); chkReal(p->retRhs, "retRhs");
  chkMac(p->fractions, "fractions"); chkMac(p->obvel, "obvel");
}

// computePressureRhs (pressure.cpp:277-299)
static void computeRhs(Solver* s, float* rhs, const Grid* vel, const Grid* flags, const mf_pressure_params* p, bool freshTemp) {
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "computePressureRhs: grid too small");
  RhsArgs a;
  a.flags = flags->di(); a.rhs = rhs; a.vel = cv3(vel);
  a.perCellCorr = p->perCellCorr ? G(p->perCellCorr)->d : nullptr;
  a.haveFractions = p->fractions ? 1 : 0; if (p->fractions) a.fractions = cv3(G(p->fractions)); else a.fractions = cv3(vel);
  a.haveObvel = p->obvel ? 1 : 0; if (p->obvel) a.obvel = cv3(G(p->obvel)); else a.obvel = cv3(vel);
  a.phi = p->phi ? G(p->phi)->d : nullptr; a.curv = p->curv ? G(p->curv)->d : nullptr;
  a.surfTens = p->surfTens; a.gfClamp = p->gfClamp;
  a.writeBorder = freshTemp ? 1 : 0;
  const bool plain = !p->fractions && !p->obvel && !(p->phi && p->curv) && !p->perCellCorr;
  if (plain && s->nx % 4 == 0) { MF_RANKS(s, r) { const Geo g = geo(s, r); MF_LAUNCH_STEP(s, k_make_rhs_plain<4>, quadTileBlocks(s, g, 4), 256, g, a.flags, rhs, a.vel, a.writeBorder, redCtx(s, r), s->redOut); } }
  else if (plain) { MF_RANKS(s, r) { const Geo g = geo(s, r); MF_LAUNCH(s, k_make_rhs_plain<1>, quadTileBlocks(s, g, 1), 256, g, a.flags, rhs, a.vel, a.writeBorder, redCtx(s, r), s->redOut); } }
  else { MF_RANKS(s, r) { const Geo g = geo(s, r); MF_LAUNCH(s, k_make_rhs, tileBlocks(s, g), 256, g, a, redCtx(s, r), s->redOut); } }
  redFinish<2>(s, RedOps<2>{ { 0, 0 } }, FinStore<2>{ s->redOut });
  if (p->enforceCompatibility) {
    fetchRed(s, 2);
    const double sum = s->redHost[0]; const int cnt = (int)s->redHost[1];
    const float corr = (float)(-sum / (float)cnt);            // rhs += (Real)(-sum / (Real)cnt)
    MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH(s, k_add_scalar, ewBlocks(s, sp.n), 256, rhs + sp.off, corr, sp.n); }
    s->rankSync();
  }
}

struct SolveTemps {   // the 7 (+4) temp grids of solvePressureSystem (pressure.cpp:332-338,412-415)
  Solver* s; float *residual, *search, *A0, *Ai, *Aj, *Ak, *tmp, *pc;
  SolveTemps(Solver* s_) : s(s_) { residual = s->getReal(); search = s->getReal(); A0 = s->getReal(); Ai = s->getReal(); Aj = s->getReal(); Ak = s->getReal(); tmp = s->getReal(); pc = nullptr; }
  ~SolveTemps() { if (pc) s->putReal(pc); s->putReal(tmp); s->putReal(Ak); s->putReal(Aj); s->putReal(Ai); s->putReal(A0); s->putReal(search); s->putReal(residual); }
};

// tmp = A search from the 16-bit row codes (+ the partial dot products): the 1-D kernel; MF_SPMV_ZM=1 selects the z-marching
// register pipeline in 3-D (bit-identical, tests compare the two; measured 50.2 us against 50.9 us at 256^3 -- the kernel is bound
// by its 229 instructions per 4-cell group, not by the re-fetches of `search` through L1, so the simpler kernel stays the default)
static void launchSpmvCode(Solver* s, const CgVecs& v, CgState* st, bool pdl) {
  if (s->is3D && std::getenv("MF_SPMV_ZM")) {
    const int zc = std::getenv("MF_SPMV_ZC") ? atoi(std::getenv("MF_SPMV_ZC")) : 32;
    MF_RANKS(s, r) { const Geo g = geo(s, r);
      const int items = s->ny * (((s->nx >> 2) + 31) >> 5), ntiles = (items + 7) / 8, chunks = (g.k1 - g.k0 + zc - 1) / zc;
      if (chunks <= 0) continue;
      const long long units = (long long)ntiles * chunks;
      const int cap = std::min(RED_MAX_BLOCKS, s->numSMs * 6);
      const int grid = (int)std::min<long long>(units, cap);
      if (pdl) { MF_LAUNCH_PDL(s, k_cg_spmv_dot_zm, grid, 256, g, v, st, redCtx(s, r), zc, ntiles, (int)units); }
      else { MF_LAUNCH(s, k_cg_spmv_dot_zm, grid, 256, g, v, st, redCtx(s, r), zc, ntiles, (int)units); } }
    return;
  }
  MF_RANKS(s, r) { const Geo g = geo(s, r); const Span sp = span(s, r);
    if (pdl) { MF_LAUNCH_PDL(s, k_cg_spmv_dot_code, redBlocks(s, sp.n, 5), 256, g, sp, v, st, redCtx(s, r)); }
    else { MF_LAUNCH(s, k_cg_spmv_dot_code, redBlocks(s, sp.n), 256, g, sp, v, st, redCtx(s, r)); } }
}

struct CgDriver {
  Solver* s; CgVecs v; CgState* st; int pcMethod;   // 0 none, 1 red-black MIC (partitioned solvers), 2 multigrid, 3 the reference's MIC(0) (mic.cuh)
  float* P; MgHierarchy* mg;
  bool v4 = false;      // nx % 4 == 0: every slab starts and ends on a 16-byte boundary -> vectorised kernels
  CgState host;
  // returns true when the preconditioner's last kernel also reduced tmp . residual and set sigma / beta (mg_apply with MgDotBeta)
  bool applyPrecond(bool init, bool wantBeta = false) {
    bool betaDone = false;
    if (pcMethod == 1) {
      const dim3 b = ijkBlockOf(s);
      MF_RANKS(s, r) { const Geo g = geo(s, r); dim3 gr(cdiv((s->nx + 1) / 2, b.x), cdiv(s->ny, b.y), cdiv(g.k1 - g.k0, b.z));
                       MF_LAUNCH(s, k_rbic_black, gr, b, g, P, v.residual, v.tmp, v.Ai, v.Aj, v.Ak, st); }
      s->rankSync();
      MF_RANKS(s, r) { const Geo g = geo(s, r); dim3 gr(cdiv((s->nx + 1) / 2, b.x), cdiv(s->ny, b.y), cdiv(g.k1 - g.k0, b.z));
                       MF_LAUNCH(s, k_rbic_red, gr, b, g, P, v.residual, v.tmp, v.Ai, v.Aj, v.Ak, st); }
      s->rankSync();
    } else if (pcMethod == 2) {
      const MgDotBeta d = { &st->sigma, &st->beta, &st->resNorm, &st->done, &st->diverged };
      betaDone = mg_apply(mg, v.tmp, v.residual, wantBeta ? &d : nullptr);
    } else if (pcMethod == 3) {
      if (v.code && s->is3D && !std::getenv("MF_MIC_GENERIC")) {
        micSweepCode(s, 1, v.code, P, v.residual, v.tmp, st);
        micSweepCode(s, 2, v.code, P, nullptr, v.tmp, st);
      } else {
        micSweep(s, 1, v.flags, v.A0, v.Ai, v.Aj, v.Ak, P, v.residual, v.tmp, st);
        micSweep(s, 2, v.flags, v.A0, v.Ai, v.Aj, v.Ak, P, nullptr, v.tmp, st);
      }
    }
    (void)init;
    return betaDone;
  }
  void fetch() { fetchWait(fetchLaunch()); }
  unsigned fetchLaunch() { return s->readSmallLaunch(s->hostInts, st, (int)sizeof(CgState) + 8); }   // + the word behind it (solveSystem: fixedDev)
  void fetchWait(unsigned ticket) {
    s->readSmallWait(ticket);
    memcpy(&host, s->hostInts, sizeof(CgState));
    if (pcMethod == 2) mg_check_errors(mg);
  }
  // doInit (conjugategrad.cpp:210-235)
  void init(float accuracy, bool useL2) {
    const long long N = s->N;
    CgState h; memset(&h, 0, sizeof(h));
    h.accuracy = accuracy; h.useL2 = useL2 ? 1 : 0; h.resNorm = 1e20f;
    s->writeSmall(st, &h, (int)sizeof(h));
    (void)N;
    slabZero(s, v.dst);
    if (v.residual != v.rhs) slabCopy(s, v.residual, v.rhs);          // (solvePressure hands its own rhs temporary over as the residual)
    const float* z = v.residual;
    if (pcMethod != 0) { applyPrecond(true); z = v.tmp; }
    // search = z in the pass that computes sigma
    MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH(s, k_cg_dot_sigma, redBlocks(s, sp.n), 256, sp.n, z + sp.off, v.residual + sp.off, v.search + sp.off, st, redCtx(s, r)); }
    redFinish<1>(s, RedOps<1>{ { 0 } }, FinSigma{ st });
    s->rankSync();
  }
  // tmp = A search, alpha (conjugategrad.cpp:246-252)
  void spmv() {
    if (v.code) launchSpmvCode(s, v, st, true);
    else { MF_RANKS(s, r) { const Geo g = geo(s, r); MF_LAUNCH_PDL(s, k_cg_spmv_dot, tileBlocks(s, g), 256, g, v, st, redCtx(s, r)); } }
    redFinish<1>(s, RedOps<1>{ { 0 } }, FinAlpha{ st });
  }
  // preconditioned iteration, first half: matrix product, x / residual update, norm and stop test (:246-277; the reference applies
  // the preconditioner before it tests the norm, but the test does not depend on it)
  template <bool V4> void headT() {
    spmv();
    MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH_PDL(s, (k_cg_update<false, V4>), redBlocks(s, sp.n), 256, sp, v, st, redCtx(s, r)); }
    redFinish<4>(s, RedOps<4>{ { 0, 1, 2, 0 } }, FinUpdate<false>{ st });
  }
  // second half: tmp = M^-1 residual, sigmaNew, beta, new search direction (:262-295)
  template <bool V4> void tailT() {
    if (!applyPrecond(false, true)) {
      MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH_PDL(s, k_cg_dot_beta<V4>, redBlocks(s, sp.n), 256, sp.n, v.tmp + sp.off, v.residual + sp.off, st, redCtx(s, r)); }
      redFinish<1>(s, RedOps<1>{ { 0 } }, FinBeta{ st });
    }
    MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH_PDL(s, k_cg_search<V4>, ewBlocks(s, sp.n / (V4 ? 4 : 1)), 256, sp.n, v.search + sp.off, v.tmp + sp.off, st); }
    s->rankSync();
  }
  template <bool V4> void iterateT() {
    if (pcMethod == 0) {
      spmv();
      MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH_PDL(s, (k_cg_update<true, V4>), redBlocks(s, sp.n), 256, sp, v, st, redCtx(s, r)); }
      redFinish<4>(s, RedOps<4>{ { 0, 1, 2, 0 } }, FinUpdate<true>{ st });
      MF_RANKS(s, r) { const Span sp = span(s, r); MF_LAUNCH_PDL(s, k_cg_search<V4>, ewBlocks(s, sp.n / (V4 ? 4 : 1)), 256, sp.n, v.search + sp.off, v.residual + sp.off, st); }
      s->rankSync();
    } else {
      headT<V4>();
      tailT<V4>();
    }
  }
  void iterate() { if (v4) iterateT<true>(); else iterateT<false>(); }
  // for (iter < maxIter) if (!iterate()) break;   host polls the device flag every `batch` iterations
  // Multigrid-preconditioned solve: the host reads the stop flag once per iteration anyway (a V-cycle is too expensive to
  // run ahead of it), so it reads it BETWEEN the two halves: the V-cycle, dot product and search update of the iteration that
  // converged -- which the reference computes into its temporary grids and never uses -- are not launched.  Same results, one
  // V-cycle less per solve.  (The GridCg object keeps whole iterations: its caller owns tmp / search and may look at them.)
  void solveSplit(int maxIter) {
    host.done = 0; host.diverged = 0; host.iterations = 0;
    bool fresh = false;
    for (int it = 0; it < maxIter; it++) {
      if (v4) headT<true>(); else headT<false>();
      MF_LAUNCH_CHECK();
      // the copy of the state is queued, then -- ahead of the decision -- the first kernel of the V-cycle the next half needs (it
      // returns at once on the device when the stop flag is set): the device works while the host waits for the flag
      const unsigned ticket = fetchLaunch();
      const bool begun = mg_apply_begin(mg, v.tmp, v.residual, &st->done);
      fetchWait(ticket); fresh = true;
      if (host.done) { if (begun) mg_apply_abort(mg); break; }
      if (v4) tailT<true>(); else tailT<false>();
      fresh = false;
    }
    if (!fresh) { MF_LAUNCH_CHECK(); fetch(); }
    if (host.diverged) mfErr("GridCg::iterate: The CG solver diverged, residual norm > 1e30, stopping.");
  }
  void solve(int maxIter, int batch) {
    int it = 0;
    host.done = 0; host.diverged = 0; host.iterations = 0;
    while (it < maxIter) {
      int nb = batch < (maxIter - it) ? batch : (maxIter - it);
      for (int b = 0; b < nb; b++) iterate();
      it += nb;
      MF_LAUNCH_CHECK();
      fetch();
      if (host.done) break;
    }
    if (maxIter <= 0) fetch();
    if (host.diverged) mfErr("GridCg::iterate: The CG solver diverged, residual norm > 1e30, stopping.");
  }
};

static CgState* cgStateBuf(Solver* s) {
  // carve the device CgState out of the reduction output buffer (16 doubles): doubles 8..15
  return reinterpret_cast<CgState*>(s->redOut + 8);
}

// solvePressureSystem (pressure.cpp:312-454)
static void solveSystem(Solver* s, float* rhs, const Grid* vel, Grid* pressure, const Grid* flags, const mf_pressure_params* p, mf_solve_info* info,
                        bool rhsScratch = false) {
  int preconditioner = p->preconditioner;
  if (!p->precondition) preconditioner = MF_PC_NONE;
  mfAssert(preconditioner >= 0 && preconditioner <= 3, "solvePressureSystem: unknown preconditioner " << preconditioner);
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "solvePressureSystem: grid too small");
  SolveTemps t(s);
  CV3 fr = p->fractions ? cv3(G(p->fractions)) : cv3(vel);
  // Row codes instead of coefficient planes: the plain Laplace matrix (no fractions, no ghost fluid) has only 0/-1 couplings and
  // an integer diagonal, and with nx % 4 == 0 the CG streams them four cells at a time.  The planes are then only written when
  // somebody reads them: the red-black IC preconditioner, or a multigrid hierarchy that has to be (re)built.
  const bool useCode = !p->fractions && !p->phi && (s->nx % 4 == 0) && !std::getenv("MF_CG_PLANES");
  const bool mgPc = (preconditioner == MF_PC_MG_STATIC || preconditioner == MF_PC_MG_DYNAMIC);
  if (mgPc && s->mg && preconditioner == MF_PC_MG_DYNAMIC) { mg_destroy(s->mg); s->mg = nullptr; }
  const bool needPlanes = !useCode || (preconditioner == MF_PC_MIC && s->is3D) || (mgPc && !(s->mg && mg_is_a_set(s->mg)));
  if (useCode && !s->cgCode) s->cgCode = (unsigned short*)s->galloc(sizeof(unsigned short) * (size_t)(s->N + 64), 1, 0, nullptr, 2);
  const float* phiD = p->phi ? G(p->phi)->d : nullptr;
  // the pinned cell's index sits right behind the CG state (redOut doubles 8..12), so that the solver's last read of the state brings
  // it to the host as well: no separate synchronisation for mf_solve_info::fixedCell
  int* fixedDev = reinterpret_cast<int*>(s->redOut + 13);
  bool fixing = (p->zeroPressureFixing || p->cgAccuracy < 1e-07f);
  // the row codes and the choice of the pinned cell depend on the flags only: reuse them while the flag grid is unchanged
  const bool codeCached = useCode && !needPlanes && !s->multi() && s->codeFlags == (const void*)flags->d && s->codeVersion == flags->version &&
                          s->codeFixing == (fixing ? 1 : 0) && !std::getenv("MF_CG_NOCODECACHE");
  s->codeFlags = nullptr;
  bool statsDone = false;
  if (codeCached) {
    if (fixing && s->codeFixedCell >= 0) CUDA_CHECK(cudaMemsetAsync(rhs + s->codeFixedCell, 0, sizeof(float), s->stream));   // fixPressure: rhs(cell) = 0
  } else
  if (useCode && needPlanes) MF_LAUNCH_IJK(s, k_make_laplace<true>, flags->di(), t.A0, t.Ai, t.Aj, t.Ak, fr, 0, phiD, p->gfClamp, s->cgCode);
  else if (useCode && fixing && !std::getenv("MF_CODE_STATS_SEPARATE")) {
    // row codes and the statistics of zeroPressureFixing in one pass over the flags
    statsDone = true;
    MF_RANKS(s, r) { const Geo g = geo(s, r);
      if (s->nx % 4 == 0) { MF_LAUNCH(s, k_make_laplace_code_stats<4>, quadTileBlocks(s, g, 4), 256, g, flags->di(), s->cgCode, redCtx(s, r), s->redOut); }
      else { MF_LAUNCH(s, k_make_laplace_code_stats<1>, quadTileBlocks(s, g, 1), 256, g, flags->di(), s->cgCode, redCtx(s, r), s->redOut); } }
  }
  else if (useCode) MF_LAUNCH_IJK_W(s, k_make_laplace_code, flags->di(), s->cgCode);
  else MF_LAUNCH_IJK(s, k_make_laplace<false>, flags->di(), t.A0, t.Ai, t.Aj, t.Ak, fr, p->fractions ? 1 : 0, phiD, p->gfClamp, (unsigned short*)nullptr);
  if (fixing && !codeCached) {
    if (statsDone) {}
    else if (s->nx % 4 == 0) { MF_RANKS(s, r) { const Geo g = geo(s, r); MF_LAUNCH(s, k_flag_stats<4>, quadTileBlocks(s, g, 4), 256, g, flags->di(), redCtx(s, r), s->redOut); } }
    else { MF_RANKS(s, r) { const Geo g = geo(s, r); MF_LAUNCH(s, k_flag_stats<1>, quadTileBlocks(s, g, 1), 256, g, flags->di(), redCtx(s, r), s->redOut); } }
    redFinish<2>(s, RedOps<2>{ { 0, 1 } }, FinStore<2>{ s->redOut });
    // every process applies the same single-cell fix (identical inputs, value 0: redundant peer writes are benign)
    if (useCode && needPlanes) MF_LAUNCH(s, (k_fix_pressure<true, true>), 1, 32, geo(s), flags->di(), s->redOut, rhs, t.A0, t.Ai, t.Aj, t.Ak, s->cgCode, fixedDev);
    else if (useCode) MF_LAUNCH(s, (k_fix_pressure<false, true>), 1, 32, geo(s), flags->di(), s->redOut, rhs, t.A0, t.Ai, t.Aj, t.Ak, s->cgCode, fixedDev);
    else MF_LAUNCH(s, (k_fix_pressure<true, false>), 1, 32, geo(s), flags->di(), s->redOut, rhs, t.A0, t.Ai, t.Aj, t.Ak, (unsigned short*)nullptr, fixedDev);
    s->rankSync();
  }
  CgDriver cg;
  cg.s = s; cg.st = cgStateBuf(s); cg.P = nullptr; cg.mg = nullptr;
  cg.v.flags = flags->di(); cg.v.dst = pressure->d; cg.v.residual = rhsScratch ? rhs : t.residual; cg.v.search = t.search; cg.v.tmp = t.tmp;
  cg.v.rhs = rhs; cg.v.A0 = t.A0; cg.v.Ai = t.Ai; cg.v.Aj = t.Aj; cg.v.Ak = t.Ak;
  cg.v.code = useCode ? s->cgCode : nullptr;
  cg.v4 = (s->nx % 4 == 0);
  int maxIter = 0, batch = 8;
  int mgLevels = 0, mgRebuilt = 0;
  int maxDim = s->nx > s->ny ? s->nx : s->ny; if (s->nz > maxDim) maxDim = s->nz;
  if (preconditioner == MF_PC_NONE || preconditioner == MF_PC_MIC) {
    maxIter = (int)(p->cgMaxIterFac * maxDim) * (s->is3D ? 1 : 4);
    // 2-D: ICP/mICP silently disabled (conjugategrad.cpp:315-321)
    if (preconditioner == MF_PC_MIC && s->is3D) {
      t.pc = s->getReal(); cg.P = t.pc;
      // single-GPU solvers run the reference's own sequential MIC(0) as a tile wavefront (bit-identical iterates); z-slab
      // partitioned solvers keep the red-black variant (a lexicographic sweep would serialise the ranks)
      if (!s->multi() && !std::getenv("MF_MIC_REDBLACK")) {
        cg.pcMethod = 3;
        micSweep(s, 0, flags->di(), t.A0, t.Ai, t.Aj, t.Ak, cg.P, nullptr, nullptr, nullptr);
      } else {
        cg.pcMethod = 1;
        maxIter *= 3;      // the red-black ordering needs about 2.5x the iterations of the sequential one (73 vs 29 on the 18x22x14 case)
        MF_LAUNCH_IJK(s, k_rbic_init, flags->di(), cg.P, t.A0, t.Ai, t.Aj, t.Ak);
      }
    } else cg.pcMethod = 0;
  } else {
    maxIter = 100; batch = 1;
    cg.pcMethod = 2;
    if (!s->mg) s->mg = mg_create(s);
    cg.mg = s->mg;
    // InitPreconditionMultigrid (conjugategrad.cpp:100-106)
    if (!mg_is_a_set(cg.mg)) { mg_set_a(cg.mg, t.A0, t.Ai, t.Aj, t.Ak); mgRebuilt = 1; }
    mg_set_coarsest_accuracy(cg.mg, p->cgAccuracy * 1E-4f);
    mg_set_smoothing(cg.mg, 1, 1);
    mgLevels = mg_num_levels(cg.mg);
  }
  cg.init(p->cgAccuracy, p->useL2Norm != 0);
  // MF_TIME_PCG: user event slots 14 / 15 bracket the iteration loop alone (bench.py: PCG time per iteration)
  const bool timeLoop = std::getenv("MF_TIME_PCG") != nullptr;
  if (timeLoop) { for (int e = 14; e < 16; e++) if (!s->userEvents[e]) CUDA_CHECK(cudaEventCreate(&s->userEvents[e])); CUDA_CHECK(cudaEventRecord(s->userEvents[14], s->stream)); }
  if (cg.pcMethod == 2 && !std::getenv("MF_CG_WHOLE_ITERATIONS")) cg.solveSplit(maxIter);
  else cg.solve(maxIter, batch);
  if (timeLoop) CUDA_CHECK(cudaEventRecord(s->userEvents[15], s->stream));
  if (info) {
    info->iterations = cg.host.iterations; info->resNorm = cg.host.resNorm; info->maxIter = maxIter;
    info->mgLevels = mgLevels; info->mgRebuilt = mgRebuilt; info->fixedCell = -1;
    if (fixing && codeCached) info->fixedCell = s->codeFixedCell;
    else if (fixing) {
      info->fixedCell = s->hostInts[sizeof(CgState) / sizeof(int)];
    }
  }
  // remember what the row codes were built from (a solve that also wrote the planes rebuilt them, which is just as valid)
  if (useCode && !s->multi() && (!fixing || codeCached || info)) {
    if (fixing && !codeCached) s->codeFixedCell = info->fixedCell;
    s->codeFlags = flags->d; s->codeVersion = flags->version; s->codeFixing = fixing ? 1 : 0;
  }
  if (cg.mg && preconditioner == MF_PC_MG_DYNAMIC) { mg_destroy(s->mg); s->mg = nullptr; }
}

static void correctVel(Solver* s, Grid* vel, const Grid* pressure, const Grid* flags, const mf_pressure_params* p) {
  MF_LAUNCH_IJK_W_PDL(s, k_correct_velocity, flags->di(), v3(vel), pressure->d);
  if (p->phi) {
    MF_LAUNCH_IJK(s, k_correct_velocity_gf, flags->di(), v3(vel), pressure->d, G(p->phi)->d, p->gfClamp,
                  p->curv ? G(p->curv)->d : nullptr, p->surfTens);
    MF_LAUNCH_IJK(s, k_replace_clamped_gf, flags->di(), v3(vel), G(p->phi)->d, p->gfClamp);
  }
}

// matrix of cgSolveDiffusion (conjugategrad.cpp:359-375): MakeLaplaceMatrix on an all-fluid dummy flag grid, then
// identity rows in obstacles and (1 + alpha*L) elsewhere
__global__ void __launch_bounds__(256) k_diffusion_matrix(Geo g, const int* __restrict__ flags, float* A0, float* Ai, float* Aj, float* Ak, float alpha) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0;
  if (i >= g.nx || j >= g.ny || k >= g.k1) return;
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;
  float a0 = 0.f, ai = 0.f, aj = 0.f, ak = 0.f;
  if (interiorCell(g, i, j, k)) { a0 = g.is3D ? 6.f : 4.f; ai = -1.f; aj = -1.f; ak = g.is3D ? -1.f : 0.f; }
  if (fIsObstacle(flags[idx])) { ai = aj = ak = 0.f; a0 = 1.f; }
  else { ai *= alpha; aj *= alpha; ak *= alpha; a0 *= alpha; a0 += 1.f; }
  A0[idx] = a0; Ai[idx] = ai; Aj[idx] = aj; Ak[idx] = ak;
}

} // namespace mf

using namespace mf;

namespace mf { void release_mg(Solver* s) { if (s->mg) { mg_destroy(s->mg); s->mg = nullptr; } }
               void release_mic(Solver* s) { if (s->mic) { micDestroy(s->mic); s->mic = nullptr; } } }

extern "C" {

void mf_pressure_params_default(mf_pressure_params* p) {
  if (!p) return;
  p->cgAccuracy = 1e-3f; p->phi = nullptr; p->perCellCorr = nullptr; p->fractions = nullptr; p->obvel = nullptr;
  p->gfClamp = 1e-04f; p->cgMaxIterFac = 1.5f; p->precondition = 1; p->preconditioner = MF_PC_MIC;
  p->enforceCompatibility = 0; p->useL2Norm = 0; p->zeroPressureFixing = 0; p->curv = nullptr; p->surfTens = 0.f; p->retRhs = nullptr;
}

int mf_compute_pressure_rhs(mf_grid* rhs, const mf_grid* vel, const mf_grid* pressure, const mf_grid* flags, const mf_pressure_params* p) {
  MF_API_BEGIN
  REQ_GRID(rhs, G(rhs)->isReal(), "rhs", "computePressureRhs");
  REQ_GRID(vel, G(vel)->isMAC(), "vel", "computePressureRhs");
  REQ_GRID(flags, G(flags)->isInt(), "flags", "computePressureRhs");
  Solver* s = G(flags)->s; s->use();
  check_same_solver(G(flags), G(vel), "computePressureRhs"); check_same_solver(G(flags), G(rhs), "computePressureRhs");
  checkParams(s, p, "computePressureRhs");
  (void)pressure;
  computeRhs(s, G(rhs)->d, G(vel), G(flags), p, false);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_solve_pressure_system(mf_grid* rhs, mf_grid* vel, mf_grid* pressure, const mf_grid* flags, const mf_pressure_params* p, mf_solve_info* info) {
  MF_API_BEGIN
  REQ_GRID(rhs, G(rhs)->isReal(), "rhs", "solvePressureSystem");
  REQ_GRID(vel, G(vel)->isMAC(), "vel", "solvePressureSystem");
  REQ_GRID(pressure, G(pressure)->isReal(), "pressure", "solvePressureSystem");
  REQ_GRID(flags, G(flags)->isInt(), "flags", "solvePressureSystem");
  Solver* s = G(flags)->s; s->use();
  check_same_solver(G(flags), G(vel), "solvePressureSystem"); check_same_solver(G(flags), G(rhs), "solvePressureSystem");
  check_same_solver(G(flags), G(pressure), "solvePressureSystem");
  checkParams(s, p, "solvePressureSystem");
  solveSystem(s, G(rhs)->d, G(vel), G(pressure), G(flags), p, info);
  MF_API_END
}

int mf_correct_velocity(mf_grid* vel, mf_grid* pressure, const mf_grid* flags, const mf_pressure_params* p) {
  MF_API_BEGIN
  REQ_GRID(vel, G(vel)->isMAC(), "vel", "correctVelocity");
  REQ_GRID(pressure, G(pressure)->isReal(), "pressure", "correctVelocity");
  REQ_GRID(flags, G(flags)->isInt(), "flags", "correctVelocity");
  Solver* s = G(flags)->s; s->use();
  check_same_solver(G(flags), G(vel), "correctVelocity"); check_same_solver(G(flags), G(pressure), "correctVelocity");
  checkParams(s, p, "correctVelocity");
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "correctVelocity: grid too small");
  correctVel(s, G(vel), G(pressure), G(flags), p);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_solve_pressure(mf_grid* vel, mf_grid* pressure, const mf_grid* flags, const mf_pressure_params* p, mf_solve_info* info) {
  MF_API_BEGIN
  REQ_GRID(vel, G(vel)->isMAC(), "vel", "solvePressure");
  REQ_GRID(pressure, G(pressure)->isReal(), "pressure", "solvePressure");
  REQ_GRID(flags, G(flags)->isInt(), "flags", "solvePressure");
  Solver* s = G(flags)->s; s->use();
  check_same_solver(G(flags), G(vel), "solvePressure"); check_same_solver(G(flags), G(pressure), "solvePressure");
  checkParams(s, p, "solvePressure");
  TmpReal rhs(s);                                    // Grid<Real> rhs(vel.getParent())  (pressure.cpp:499)
  computeRhs(s, rhs.p, G(vel), G(flags), p, true);
  solveSystem(s, rhs.p, G(vel), G(pressure), G(flags), p, info, /* rhs is not needed afterwards */ !p->retRhs);
  correctVel(s, G(vel), G(pressure), G(flags), p);
  if (p->retRhs) slabCopy(s, G(p->retRhs)->d, rhs.p);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_release_mg(mf_solver* sp) {
  MF_API_BEGIN
  mfAssert(sp, "releaseMG: null solver");
  S(sp)->use();
  release_mg(S(sp));
  MF_API_END
}

int mf_make_laplace_matrix(const mf_grid* flags, mf_grid* A0, mf_grid* Ai, mf_grid* Aj, mf_grid* Ak, const mf_grid* fractions) {
  MF_API_BEGIN
  REQ_GRID(flags, G(flags)->isInt(), "flags", "MakeLaplaceMatrix");
  REQ_GRID(A0, G(A0)->isReal(), "A0", "MakeLaplaceMatrix"); REQ_GRID(Ai, G(Ai)->isReal(), "Ai", "MakeLaplaceMatrix");
  REQ_GRID(Aj, G(Aj)->isReal(), "Aj", "MakeLaplaceMatrix"); REQ_GRID(Ak, G(Ak)->isReal(), "Ak", "MakeLaplaceMatrix");
  if (fractions) REQ_GRID(fractions, G(fractions)->isMAC(), "fractions", "MakeLaplaceMatrix");
  Solver* s = G(flags)->s; s->use();
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "MakeLaplaceMatrix: grid too small");
  CV3 fr; fr.x = fr.y = fr.z = nullptr; if (fractions) fr = cv3(G(fractions));
  MF_LAUNCH_IJK(s, k_make_laplace<false>, G(flags)->di(), G(A0)->d, G(Ai)->d, G(Aj)->d, G(Ak)->d, fr, fractions ? 1 : 0, (const float*)nullptr, 0.f, (unsigned short*)nullptr);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_apply_matrix(const mf_grid* flags, mf_grid* dst, const mf_grid* src, const mf_grid* A0, const mf_grid* Ai, const mf_grid* Aj, const mf_grid* Ak) {
  MF_API_BEGIN
  REQ_GRID(flags, G(flags)->isInt(), "flags", "ApplyMatrix");
  REQ_GRID(dst, G(dst)->isReal(), "dst", "ApplyMatrix"); REQ_GRID(src, G(src)->isReal(), "src", "ApplyMatrix");
  REQ_GRID(A0, G(A0)->isReal(), "A0", "ApplyMatrix"); REQ_GRID(Ai, G(Ai)->isReal(), "Ai", "ApplyMatrix");
  REQ_GRID(Aj, G(Aj)->isReal(), "Aj", "ApplyMatrix"); REQ_GRID(Ak, G(Ak)->isReal(), "Ak", "ApplyMatrix");
  mfAssert(G(dst)->d != G(src)->d, "ApplyMatrix: dst and src must differ");
  Solver* s = G(flags)->s; s->use();
  MF_LAUNCH_IJK(s, k_apply_matrix, G(flags)->di(), G(dst)->d, G(src)->d, G(A0)->d, G(Ai)->d, G(Aj)->d, G(Ak)->d);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_apply_laplace_rowcode(const mf_grid* flags, mf_grid* dst, const mf_grid* src) {
  MF_API_BEGIN
  REQ_GRID(flags, G(flags)->isInt(), "flags", "mf_apply_laplace_rowcode");
  REQ_GRID(dst, G(dst)->isReal(), "dst", "mf_apply_laplace_rowcode"); REQ_GRID(src, G(src)->isReal(), "src", "mf_apply_laplace_rowcode");
  mfAssert(G(dst)->d != G(src)->d, "mf_apply_laplace_rowcode: dst and src must differ");
  Solver* s = G(flags)->s; s->use();
  mfAssert(s->nx % 4 == 0, "mf_apply_laplace_rowcode: nx must be a multiple of 4");
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "mf_apply_laplace_rowcode: grid too small");
  if (!s->cgCode) s->cgCode = (unsigned short*)s->galloc(sizeof(unsigned short) * (size_t)(s->N + 64), 1, 0, nullptr, 2);
  s->codeFlags = nullptr;       // the solver's cached row codes are overwritten here
  MF_LAUNCH_IJK_W(s, k_make_laplace_code, G(flags)->di(), s->cgCode);
  CgState* st = cgStateBuf(s);
  CUDA_CHECK(cudaMemsetAsync(st, 0, sizeof(CgState), s->stream));
  CgVecs v; memset(&v, 0, sizeof(v));
  v.search = G(src)->d; v.tmp = G(dst)->d; v.code = s->cgCode;
  launchSpmvCode(s, v, st, false);
  redFinish<1>(s, RedOps<1>{ { 0 } }, FinAlpha{ st });
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_cg_solve_diffusion(const mf_grid* flagsp, mf_grid* gridp, float alpha, float cgMaxIterFac, float cgAccuracy) {
  MF_API_BEGIN
  REQ_GRID(flagsp, G(flagsp)->isInt(), "flags", "cgSolveDiffusion");
  mfAssert(gridp, "cgSolveDiffusion: argument 'grid' is missing");
  const Grid* flags = G(flagsp); Grid* grid = G(gridp); Solver* s = flags->s; s->use();
  check_same_solver(flags, grid, "cgSolveDiffusion");
  mfAssert(grid->isReal() || grid->isVec(), "cgSolveDiffusion: Grid Type is not supported (only Real, Vec3, MAC, or Levelset)");
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "cgSolveDiffusion: grid too small");
  SolveTemps t(s);
  TmpReal rhs(s);
  MF_LAUNCH_IJK(s, k_diffusion_matrix, flags->di(), t.A0, t.Ai, t.Aj, t.Ak, alpha);
  int maxDim = s->nx > s->ny ? s->nx : s->ny; if (s->nz > maxDim) maxDim = s->nz;
  const int maxIter = (int)(cgMaxIterFac * maxDim) * (s->is3D ? 1 : 4);
  const int ncomp = grid->isVec() ? (s->is3D ? 3 : 2) : 1;
  for (int c = 0; c < ncomp; c++) {
    float* u = grid->isVec() ? grid->comp(c) : grid->d;
    slabCopy(s, rhs.p, u);
    CgDriver cg;
    cg.s = s; cg.st = cgStateBuf(s); cg.P = nullptr; cg.mg = nullptr; cg.pcMethod = 0;
    cg.v.flags = flags->di(); cg.v.dst = u; cg.v.residual = t.residual; cg.v.search = t.search; cg.v.tmp = t.tmp;
    cg.v.rhs = rhs.p; cg.v.A0 = t.A0; cg.v.Ai = t.Ai; cg.v.Aj = t.Aj; cg.v.Ak = t.Ak;
    cg.v.code = nullptr; cg.v4 = (s->nx % 4 == 0);
    cg.init(cgAccuracy, true);      // GridCgInterface() defaults to the L2 norm (conjugategrad.h:31)
    cg.solve(maxIter, 8);
  }
  MF_API_END
}

// ---- GridCg as an object (GridCgInterface, conjugategrad.h:27-58) ----------------------------------------------------------
namespace mf {
struct CgObject {
  Solver* s;
  CgDriver drv;
  CgState* state;       // this object's device state (several GridCg objects may be alive on one solver)
  const Grid *A0, *Ai, *Aj, *Ak; const Grid* flags;
  float accuracy; bool useL2; bool inited; int method; bool rebuildMg;
  float* pc;            // red-black IC diagonal (from the solver's pool while a preconditioned solve is set up)
};
static void cgObjInit(CgObject* c) {
  Solver* s = c->s;
  CgDriver& d = c->drv;
  d.pcMethod = 0; d.P = nullptr; d.mg = nullptr;
  if ((c->method == 1 || c->method == 2) && s->is3D) {
    if (!c->pc) c->pc = s->getReal();
    d.P = c->pc;
    if (c->method == 2 && !s->multi() && !std::getenv("MF_MIC_REDBLACK")) {    // PC_mICP: the reference's MIC(0), bit for bit (mic.cuh)
      d.pcMethod = 3;
      micSweep(s, 0, c->flags->di(), c->A0->d, c->Ai->d, c->Aj->d, c->Ak->d, d.P, nullptr, nullptr, nullptr);
    } else {
      d.pcMethod = 1;
      MF_LAUNCH_IJK(s, k_rbic_init, c->flags->di(), d.P, c->A0->d, c->Ai->d, c->Aj->d, c->Ak->d);
    }
  } else if (c->method == 3) {
    if (c->rebuildMg && s->mg) { mg_destroy(s->mg); s->mg = nullptr; }
    c->rebuildMg = false;
    if (!s->mg) s->mg = mg_create(s);
    d.pcMethod = 2; d.mg = s->mg;
    if (!mg_is_a_set(d.mg)) mg_set_a(d.mg, c->A0->d, c->Ai->d, c->Aj->d, c->Ak->d);      // InitPreconditionMultigrid (conjugategrad.cpp:100-106)
    mg_set_coarsest_accuracy(d.mg, c->accuracy * 1E-4f);
    mg_set_smoothing(d.mg, 1, 1);
  }
  d.init(c->accuracy, c->useL2);
  c->inited = true;
}
}

mf_cg* mf_cg_create(mf_grid* dst, mf_grid* rhs, mf_grid* residual, mf_grid* search, const mf_grid* flags, mf_grid* tmp,
                    mf_grid* A0, mf_grid* Ai, mf_grid* Aj, mf_grid* Ak) {
  mf_cg* out = nullptr;
  auto body = [&]() -> int {
    MF_API_BEGIN
    REQ_GRID(flags, G(flags)->isInt(), "flags", "GridCg");
    const mf_grid* gs[9] = { dst, rhs, residual, search, tmp, A0, Ai, Aj, Ak };
    const char* names[9] = { "dst", "rhs", "residual", "search", "tmp", "A0", "Ai", "Aj", "Ak" };
    for (int i = 0; i < 9; i++) {
      mfAssert(gs[i] && G(gs[i])->isReal(), "GridCg: argument '" << names[i] << "' has the wrong grid type or is missing");
      check_same_solver(G(flags), G(gs[i]), "GridCg");
    }
    Solver* s = G(flags)->s; s->use();
    mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "GridCg: grid too small");
    CgObject* c = new CgObject();
    c->s = s; c->A0 = G(A0); c->Ai = G(Ai); c->Aj = G(Aj); c->Ak = G(Ak); c->flags = G(flags);
    c->accuracy = 1e-6f; c->useL2 = true; c->inited = false; c->method = 0; c->rebuildMg = false; c->pc = nullptr;
    CUDA_CHECK(cudaMalloc(&c->state, sizeof(CgState) + 8));      // (+ the word CgDriver::fetchLaunch copies with the state)
    CUDA_CHECK(cudaMemset(c->state, 0, sizeof(CgState) + 8));
    CgDriver& d = c->drv;
    d.s = s; d.st = c->state; d.P = nullptr; d.mg = nullptr; d.pcMethod = 0;
    d.v.flags = G(flags)->di(); d.v.dst = G(dst)->d; d.v.residual = G(residual)->d; d.v.search = G(search)->d; d.v.tmp = G(tmp)->d;
    d.v.rhs = G(rhs)->d; d.v.A0 = G(A0)->d; d.v.Ai = G(Ai)->d; d.v.Aj = G(Aj)->d; d.v.Ak = G(Ak)->d;
    d.v.code = nullptr; d.v4 = (s->nx % 4 == 0);
    memset(&d.host, 0, sizeof(d.host)); d.host.resNorm = 1e20f;
    out = reinterpret_cast<mf_cg*>(c);
    MF_API_END
  };
  return body() == 0 ? out : nullptr;
}
void mf_cg_destroy(mf_cg* cg) {
  if (!cg) return;
  CgObject* c = reinterpret_cast<CgObject*>(cg);
  cudaStreamSynchronize(c->s->stream);
  if (c->pc) c->s->putReal(c->pc);
  cudaFree(c->state);
  delete c;
}
#define CG_OBJ(fn) mfAssert(cg, fn ": null GridCg"); CgObject* c = reinterpret_cast<CgObject*>(cg); c->s->use();
int mf_cg_set_ic_preconditioner(mf_cg* cg, int method) {
  MF_API_BEGIN
  CG_OBJ("setICPreconditioner")
  mfAssert(method == 1 || method == 2, "GridCg<APPLYMAT>::setICPreconditioner: Invalid method specified.");
  c->method = c->s->is3D ? method : 0;      // 2-D: disabled (conjugategrad.cpp:315-321)
  MF_API_END
}
int mf_cg_set_mg_preconditioner(mf_cg* cg, int method, int rebuild) {
  MF_API_BEGIN
  CG_OBJ("setMGPreconditioner")
  mfAssert(method == 3, "GridCg<APPLYMAT>::setMGPreconditioner: Invalid method specified.");
  c->method = 3; c->rebuildMg = rebuild != 0;
  MF_API_END
}
int mf_cg_set_accuracy(mf_cg* cg, float accuracy) {
  MF_API_BEGIN
  CG_OBJ("setAccuracy")
  c->accuracy = accuracy;
  if (c->inited) {        // takes effect on the next iterate(), as mAccuracy does
    c->drv.fetch();
    CgState h = c->drv.host; h.accuracy = accuracy;
    c->s->writeSmall(c->state, &h, (int)sizeof(h));
  }
  MF_API_END
}
int mf_cg_set_use_l2_norm(mf_cg* cg, int useL2) {
  MF_API_BEGIN
  CG_OBJ("setUseL2Norm")
  c->useL2 = useL2 != 0;
  if (c->inited) {
    c->drv.fetch();
    CgState h = c->drv.host; h.useL2 = useL2 ? 1 : 0;
    c->s->writeSmall(c->state, &h, (int)sizeof(h));
  }
  MF_API_END
}
int mf_cg_force_reinit(mf_cg* cg) {
  MF_API_BEGIN
  CG_OBJ("forceReinit")
  c->inited = false;
  MF_API_END
}
int mf_cg_iterate(mf_cg* cg, int* notConverged) {
  MF_API_BEGIN
  CG_OBJ("iterate")
  if (!c->inited) cgObjInit(c);
  c->drv.iterate();
  MF_LAUNCH_CHECK();
  c->drv.fetch();
  if (c->drv.host.diverged) mfErr("GridCg::iterate: The CG solver diverged, residual norm > 1e30, stopping.");
  if (notConverged) *notConverged = c->drv.host.done ? 0 : 1;
  MF_API_END
}
int mf_cg_solve(mf_cg* cg, int maxIter) {
  MF_API_BEGIN
  CG_OBJ("solve")
  if (maxIter > 0) {
    if (!c->inited) cgObjInit(c);
    c->drv.solve(maxIter, c->drv.pcMethod == 2 ? 1 : 8);
  }
  MF_API_END
}
int mf_cg_get(mf_cg* cg, float* sigma, float* iterations, float* resNorm, float* accuracy) {
  MF_API_BEGIN
  CG_OBJ("get")
  if (sigma) *sigma = c->drv.host.sigma;
  if (iterations) *iterations = (float)c->drv.host.iterations;
  if (resNorm) *resNorm = c->drv.host.resNorm;
  if (accuracy) *accuracy = c->accuracy;
  MF_API_END
}

int mf_mg_info(mf_solver* sp, int* levels, int* usedHostFallback) {
  MF_API_BEGIN
  mfAssert(sp, "mf_mg_info: null solver");
  Solver* s = S(sp);
  if (levels) *levels = s->mg ? mg_num_levels(s->mg) : 0;
  if (usedHostFallback) *usedHostFallback = s->mg ? mg_used_host_fallback(s->mg) : 0;
  MF_API_END
}
int mf_mg_level_size(mf_solver* sp, int l, int* size3) {
  MF_API_BEGIN
  mfAssert(sp && size3, "mf_mg_level_size: null argument");
  Solver* s = S(sp);
  mfAssert(s->mg, "mf_mg_level_size: no multigrid hierarchy cached on this solver");
  mfAssert(mg_level_size(s->mg, l, &size3[0], &size3[1], &size3[2]) > 0, "mf_mg_level_size: no such level");
  MF_API_END
}
int mf_mg_level_format(mf_solver* sp, int l, int* bytesPerEntry) {
  MF_API_BEGIN
  mfAssert(sp && bytesPerEntry, "mf_mg_level_format: null argument");
  Solver* s = S(sp);
  mfAssert(s->mg && l >= 0 && l < mg_num_levels(s->mg), "mf_mg_level_format: no such level");
  *bytesPerEntry = mg_level_format(s->mg, l);
  MF_API_END
}
int mf_mg_download_level(mf_solver* sp, int l, void* type, void* A) {
  MF_API_BEGIN
  mfAssert(sp, "mf_mg_download_level: null solver");
  Solver* s = S(sp); s->use();
  mfAssert(s->mg && l >= 0 && l < mg_num_levels(s->mg), "mf_mg_download_level: no such level");
  mg_download_level(s->mg, l, (unsigned char*)type, (float*)A);
  MF_API_END
}
int mf_mg_vcycle(mf_solver* sp, const mf_grid* A0, const mf_grid* Ai, const mf_grid* Aj, const mf_grid* Ak,
                 const mf_grid* rhs, mf_grid* dst, float coarsestAccuracy) {
  MF_API_BEGIN
  mfAssert(sp, "mf_mg_vcycle: null solver");
  Solver* s = S(sp); s->use();
  REQ_GRID(rhs, G(rhs)->isReal(), "rhs", "mf_mg_vcycle"); REQ_GRID(dst, G(dst)->isReal(), "dst", "mf_mg_vcycle");
  if (A0) {
    REQ_GRID(A0, G(A0)->isReal(), "A0", "mf_mg_vcycle"); REQ_GRID(Ai, G(Ai)->isReal(), "Ai", "mf_mg_vcycle");
    REQ_GRID(Aj, G(Aj)->isReal(), "Aj", "mf_mg_vcycle"); REQ_GRID(Ak, G(Ak)->isReal(), "Ak", "mf_mg_vcycle");
    release_mg(s);
    s->mg = mg_create(s);
    mg_set_a(s->mg, G(A0)->d, G(Ai)->d, G(Aj)->d, G(Ak)->d);
    mg_set_smoothing(s->mg, 1, 1);
  } else mfAssert(s->mg && mg_is_a_set(s->mg), "mf_mg_vcycle: no cached hierarchy to re-apply (A0 == NULL)");
  mg_set_coarsest_accuracy(s->mg, coarsestAccuracy);
  mg_apply(s->mg, G(dst)->d, G(rhs)->d);
  MF_LAUNCH_CHECK();
  MF_API_END
}

} // extern "C"
