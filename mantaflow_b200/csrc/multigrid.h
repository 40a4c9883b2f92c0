// multigrid.h -- internal interface of the device GridMg (reference: source/multigrid.h:31-137).
#pragma once
#include "common.cuh"

namespace mf {

struct MgHierarchy;

MgHierarchy* mg_create(Solver* s);                         // GridMg::GridMg (multigrid.cpp:220)
void mg_destroy(MgHierarchy* mg);
bool mg_is_a_set(const MgHierarchy* mg);                   // GridMg::isASet
void mg_set_a(MgHierarchy* mg, const float* A0, const float* Ai, const float* Aj, const float* Ak);  // GridMg::setA (:386)
void mg_set_coarsest_accuracy(MgHierarchy* mg, float acc); // setCoarsestLevelAccuracy
void mg_set_smoothing(MgHierarchy* mg, int pre, int post); // setSmoothing
int  mg_num_levels(const MgHierarchy* mg);
// ApplyPreconditionMultigrid (conjugategrad.cpp:162-167): dst = one V-cycle on A*dst = rhs from dst = 0
// dot != nullptr: the caller wants sigmaNew = dst . rhs and beta = sigmaNew / sigma next (GridCg::iterate, conjugategrad.cpp:279-295).
// When the V-cycle ends in the level-0 wavefront up leg, that kernel reduces the dot product while it stores dst (its last block
// runs the update of the CG scalars) and mg_apply returns true; otherwise false and the caller launches its own reduction.
struct MgDotBeta { float* sigma; float* beta; const float* resNorm; int* done; int* diverged; };
bool mg_apply(MgHierarchy* mg, float* dst, const float* rhs, const MgDotBeta* dot = nullptr);
// The PCG driver reads its stop flag on the host between the two halves of an iteration.  mg_apply_begin queues the first kernel of
// the V-cycle mg_apply(mg, dst, rhs) would run (the level-0 down leg; it returns at once when *skip != 0 on the device) BEFORE the host
// waits for the flag, so the device is busy while the host decides; then either mg_apply(mg, dst, rhs, ...) continues the cycle or
// mg_apply_abort drops it.  Returns false (nothing queued) when the cycle does not start with that kernel.
bool mg_apply_begin(MgHierarchy* mg, float* dst, const float* rhs, const int* skip);
void mg_apply_abort(MgHierarchy* mg);
void mg_check_errors(const MgHierarchy* mg);               // raises if a device-side barrier of the V-cycle timed out (call after a stream sync)
// test hooks
int  mg_level_size(const MgHierarchy* mg, int l, int* sx, int* sy, int* sz);
void mg_download_level(MgHierarchy* mg, int l, unsigned char* type, float* A);  // A: n*stencil floats in the reference's interleaved order
int  mg_used_host_fallback(const MgHierarchy* mg);
int  mg_level_format(const MgHierarchy* mg, int l);   // bytes per stored operator entry

} // namespace mf
