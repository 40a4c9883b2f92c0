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
void mg_check_errors(const MgHierarchy* mg);               // raises if a device-side barrier of the V-cycle timed out (call after a stream sync)
// test hooks
int  mg_level_size(const MgHierarchy* mg, int l, int* sx, int* sy, int* sz);
void mg_download_level(MgHierarchy* mg, int l, unsigned char* type, float* A);  // A: n*stencil floats in the reference's interleaved order
int  mg_used_host_fallback(const MgHierarchy* mg);
int  mg_level_format(const MgHierarchy* mg, int l);   // bytes per stored operator entry

} // namespace mf
