// shapes.cu -- scene-setup plugins that run inside the step loop of the reference's scenes, SURVEY 8(f) rank 1:
// Shape::applyToGrid / computeLevelset for Box, Sphere, Cylinder (source/shapes.cpp: ApplyShapeToGrid :40, ApplyShapeToMACGrid :62,
// Box::isInside :151, BoxSDF :178, Sphere::isInside :240, SphereSDF :309, Cylinder::isInside :324, CylinderSDF :367),
// FlagGrid::updateFromLevelset (grid.cpp:910), LevelsetGrid::join (levelset.cpp:108), updateFractions / setObstacleFlags
// (plugin/initplugins.cpp:437, :473).  Float expressions keep the reference's order (the library is built with -fmad=false).
#include "common.cuh"
#include "sample.cuh"
#include "shape_dev.cuh"
#include <cmath>

namespace mf {

#define IJK_ALL \
  const int i = blockIdx.x * blockDim.x + threadIdx.x; \
  const int j = blockIdx.y * blockDim.y + threadIdx.y; \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0; \
  if (i >= g.nx || j >= g.ny || k >= g.k1) return; \
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;

// ApplyShapeToGrid<T>: NC components of a cell-centred grid (int / Real / Vec3)
// The apply kernels visit only a conservative bounding box of the shape (shapeBounds below): Geo carries the box's z range,
// i0 / j0 shift the launch in x / y.  Every cell the reference's full-grid loop would change lies inside the box.
#define IJK_BOX \
  const int i = blockIdx.x * blockDim.x + threadIdx.x + i0; \
  const int j = blockIdx.y * blockDim.y + threadIdx.y + j0; \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0; \
  if (i >= g.nx || j >= g.ny || k >= g.k1) return; \
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;

template <class T, int NC>
__global__ void __launch_bounds__(256) k_shape_apply(Geo g, int i0, int j0, ShapeD s, T* __restrict__ d, long long np, T v0, T v1, T v2, const int* __restrict__ respect) {
  IJK_BOX
  if (respect && fIsObstacle(respect[idx])) return;
  if (!shapeInside(s, i + 0.5f, j + 0.5f, k + 0.5f)) return;
  d[idx] = v0;
  if (NC == 3) { d[idx + np] = v1; d[idx + 2 * np] = v2; }
}
// ApplyShapeToMACGrid: each face tested at its own position
__global__ void __launch_bounds__(256) k_shape_apply_mac(Geo g, int i0, int j0, ShapeD s, V3 d, float vx, float vy, float vz, const int* __restrict__ respect) {
  IJK_BOX
  if (respect && fIsObstacle(respect[idx])) return;
  if (shapeInside(s, (float)i, j + 0.5f, k + 0.5f)) d.x[idx] = vx;
  if (shapeInside(s, i + 0.5f, (float)j, k + 0.5f)) d.y[idx] = vy;
  if (shapeInside(s, i + 0.5f, j + 0.5f, (float)k)) d.z[idx] = vz;
}
// BoxSDF / SphereSDF / CylinderSDF
__global__ void __launch_bounds__(256) k_shape_sdf(Geo g, ShapeD s, float* __restrict__ phi) {
  IJK_ALL
  phi[idx] = shapeSdf(g, s, i, j, k);
}

// FlagGrid::updateFromLevelset (grid.cpp:910-920); invalidTimeValue() = -1000 (fastmarch.h:134)
__global__ void __launch_bounds__(256) k_flags_from_levelset(Geo g, int* __restrict__ flags, const float* __restrict__ phi) {
  IJK_ALL
  int f = flags[idx];
  if (fIsObstacle(f) || fIsOutflow(f)) return;
  const float p = phi[idx];
  if (p <= -1000.f) return;
  f &= ~(MF_EMPTY | MF_FLUID);
  f |= (p <= 0.f) ? MF_FLUID : MF_EMPTY;
  flags[idx] = f;
}
// KnJoin (levelset.cpp:108): a = min(a, b)
__global__ void __launch_bounds__(256) k_ls_join(Geo g, float* __restrict__ a, const float* __restrict__ b) {
  IJK_ALL
  a[idx] = fmin2(a[idx], b[idx]);
}

// calcFraction (initplugins.cpp:356-370)
__device__ __forceinline__ float calcFraction(float phi1, float phi2, float fracThreshold) {
  if (phi1 > 0.f && phi2 > 0.f) return 1.f;
  if (phi1 < 0.f && phi2 < 0.f) return 0.f;
  if (phi2 < phi1) { const float t = phi1; phi1 = phi2; phi2 = t; }
  const float denom = phi1 - phi2;
  if ((double)denom > -1e-04) return 0.5f;
  float frac = (float)(1. - (double)(phi1 / denom));
  if (frac < fracThreshold) frac = 0.f;
  return fmin2(1.f, frac);
}
__device__ __forceinline__ bool fOpenish(int f) { return (f & (MF_INFLOW | MF_OUTFLOW | MF_OPEN)) != 0; }
// KnUpdateFractions (initplugins.cpp:372-434), restated for the reference's SERIAL ijk order: a cell's "max side" writes go to
// the next cell along that axis, which is visited later and overwrites them unless it lies on the domain border (bnd = 1 skips
// it) -- so only border targets keep them.  The cell's own values: calcFraction, then 1 if a "min side" condition holds.
__global__ void __launch_bounds__(256) k_update_fractions(Geo g, const int* __restrict__ flags, const float* __restrict__ phi, V3 fr, int w, float thr) {
  IJK_ALL
  const long long Y = g.nx, Z = g.Z;
  const bool interior = i >= 1 && j >= 1 && i < g.nx - 1 && j < g.ny - 1 && (!g.is3D || (k >= 1 && k < g.nz - 1));
  if (interior) {
    const float p = phi[idx];
    float fx = calcFraction(p, phi[idx - 1], thr), fy = calcFraction(p, phi[idx - Y], thr), fz = g.is3D ? calcFraction(p, phi[idx - Z], thr) : 0.f;
    if (!(p < 0.f)) {
      bool one = false;
      if (i <= w + 1 && fOpenish(flags[idx - 1])) one = true;
      if (j <= w + 1 && fOpenish(flags[idx - Y])) one = true;
      if (g.is3D && k <= w + 1 && fOpenish(flags[idx - Z])) one = true;
      if (one) { fx = 1.f; fy = 1.f; if (g.is3D) fz = 1.f; }
    }
    fr.x[idx] = fx; fr.y[idx] = fy; fr.z[idx] = fz;
    return;
  }
  // border cell: 0 (setConst), or 1 when the interior neighbour below it along an axis has a "max side" hit
  float v = 0.f;
  auto srcOk = [&](int si, int sj, int sk) {   // the writing cell must be interior, outside the obstacle, and satisfy its range test
    const bool in = si >= 1 && sj >= 1 && si < g.nx - 1 && sj < g.ny - 1 && (!g.is3D || (sk >= 1 && sk < g.nz - 1));
    if (!in) return false;
    return !(phi[(long long)si + Y * sj + (long long)g.nx * g.ny * sk] < 0.f);
  };
  const bool openHere = fOpenish(flags[idx]);
  if (openHere) {
    if (i >= 1 && srcOk(i - 1, j, k) && (i - 1) >= g.nx - w - 2) v = 1.f;                 // max x: source (i-1,j,k) tests flags(i,j,k)
    if (j >= 1 && srcOk(i, j - 1, k) && (j - 1) >= g.ny - w - 2) v = 1.f;                 // max y
    if (g.is3D && k >= 1 && srcOk(i, j, k - 1) && j >= g.nz - w - 2) v = 1.f;             // max z (sic: the reference tests j, :423)
  }
  fr.x[idx] = v; fr.y[idx] = v; fr.z[idx] = g.is3D ? v : 0.f;
}
// KnUpdateFlagsObs (initplugins.cpp:441-468), bnd = boundaryWidth
__global__ void __launch_bounds__(256) k_set_obstacle_flags(Geo g, int* __restrict__ flags, CV3 fr, int haveFr, const float* __restrict__ phiObs,
                                                            const float* __restrict__ phiOut, const float* __restrict__ phiIn, int bnd) {
  IJK_ALL
  if (i < bnd || j < bnd || i >= g.nx - bnd || j >= g.ny - bnd || (g.is3D && (k < bnd || k >= g.nz - bnd))) return;
  bool isObs = false;
  if (haveFr) {
    float f = 0.f;
    f += fr.x[idx]; f += fr.x[idx + 1]; f += fr.y[idx]; f += fr.y[idx + g.nx];
    if (g.is3D) { f += fr.z[idx]; f += fr.z[idx + g.Z]; }
    if (f == 0.f) isObs = true;
  } else if (phiObs[idx] < 0.f) isObs = true;
  const bool isOut = phiOut && phiOut[idx] < 0.f, isIn = phiIn && phiIn[idx] < 0.f;
  int f;
  if (isObs) f = MF_OBSTACLE;
  else if (isIn) f = MF_FLUID | MF_INFLOW;
  else if (isOut) f = MF_EMPTY | MF_OUTFLOW;
  else f = MF_EMPTY;
  flags[idx] = f;
}

} // namespace mf

using namespace mf;

// Cell index box [lo, hi) that contains every cell / face position p with shapeInside(p) (positions are i, i + 0.5): the box
// itself, the sphere's scaled radius, the cylinder's bounding sphere -- each widened by two cells.  Non-finite or huge
// parameters fall back to the whole grid.
struct ShapeBox { int lo[3], hi[3]; };
static ShapeBox shapeBounds(const Solver* s, const ShapeD& d) {
  const int n[3] = {s->nx, s->ny, s->nz};
  ShapeBox b;
  for (int c = 0; c < 3; c++) {
    double lo, hi;
    if (d.type == MF_SHAPE_BOX) { lo = d.a[c]; hi = d.b[c]; }
    else if (d.type == MF_SHAPE_SPHERE) { const double e = std::fabs((double)d.r * d.b[c]); lo = d.a[c] - e; hi = d.a[c] + e; }
    else { const double e = std::sqrt((double)d.zlen * d.zlen + (double)d.r * d.r); lo = d.a[c] - e; hi = d.a[c] + e; }
    b.lo[c] = 0; b.hi[c] = n[c];
    if (std::isfinite(lo) && std::isfinite(hi) && std::fabs(lo) < 1e9 && std::fabs(hi) < 1e9) {
      b.lo[c] = std::max(0, (int)std::floor(lo) - 2);
      b.hi[c] = std::min(n[c], (int)std::ceil(hi) + 3);
    }
  }
  if (!s->is3D) { b.lo[2] = 0; b.hi[2] = 1; }
  b.lo[0] &= ~31;                                   // whole 128-byte lines in x
  return b;
}
// launch an apply kernel (Geo, i0, j0, ...) over the part of the box that lies in each local rank's slab
#define MF_LAUNCH_BOX(s, box, kern, ...) do { const dim3 _b = mf::ijkBlockOf(s); MF_RANKS(s, _r) { mf::Geo _g = mf::geo((s), _r); \
  _g.k0 = std::max(_g.k0, (box).lo[2]); _g.k1 = std::min(_g.k1, (box).hi[2]); \
  if (_g.k1 > _g.k0 && (box).hi[0] > (box).lo[0] && (box).hi[1] > (box).lo[1]) { \
    const dim3 _gr(cdiv((box).hi[0] - (box).lo[0], _b.x), cdiv((box).hi[1] - (box).lo[1], _b.y), cdiv(_g.k1 - _g.k0, _b.z)); \
    mf::Geo _gg = _g; _gg.nx = (s)->nx; _gg.ny = (s)->ny; \
    MF_LAUNCH(s, kern, _gr, _b, _gg, (box).lo[0], (box).lo[1], __VA_ARGS__); } } (s)->rankSync(); } while (0)

extern "C" {

int mf_shape_apply(const mf_shape* sh, mf_grid* gp, float vx, float vy, float vz, int ivalue, const mf_grid* respectFlags) {
  MF_API_BEGIN
  mfAssert(gp, "Shape::applyToGrid: argument 'grid' is missing");
  if (respectFlags) REQ_GRID(respectFlags, G(respectFlags)->isInt(), "respectFlags", "Shape::applyToGrid");
  Grid* gr = G(gp); Solver* s = gr->s; s->use();
  const ShapeD d = shapeOf(sh);
  const int* rf = respectFlags ? G(respectFlags)->di() : nullptr;
  const ShapeBox box = shapeBounds(s, d);
  if (gr->isInt()) { markFlagsChanged(gr); MF_LAUNCH_BOX(s, box, (k_shape_apply<int, 1>), d, gr->di(), 0LL, ivalue, 0, 0, rf); }
  else if (gr->isReal()) MF_LAUNCH_BOX(s, box, (k_shape_apply<float, 1>), d, gr->d, 0LL, vx, 0.f, 0.f, rf);
  else if (gr->isMAC()) MF_LAUNCH_BOX(s, box, k_shape_apply_mac, d, v3(gr), vx, vy, vz, rf);
  else if (gr->isVec()) MF_LAUNCH_BOX(s, box, (k_shape_apply<float, 3>), d, gr->d, s->Np, vx, vy, vz, rf);
  else mfErr("Shape::applyToGrid(): unknown grid type");
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_shape_levelset(const mf_shape* sh, mf_grid* phi) {
  MF_API_BEGIN
  REQ_GRID(phi, G(phi)->isReal(), "phi", "Shape::computeLevelset");
  Grid* gr = G(phi); Solver* s = gr->s; s->use();
  const ShapeD d = shapeOf(sh);
  MF_LAUNCH_IJK(s, k_shape_sdf, d, gr->d);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_flags_update_from_levelset(mf_grid* flags, const mf_grid* phi) {
  MF_API_BEGIN
  REQ_GRID(flags, G(flags)->isInt(), "flags", "FlagGrid::updateFromLevelset");
  REQ_GRID(phi, G(phi)->isReal(), "levelset", "FlagGrid::updateFromLevelset");
  Grid* f = G(flags); Solver* s = f->s; s->use();
  check_same_solver(f, G(phi), "FlagGrid::updateFromLevelset");
  markFlagsChanged(f);
  MF_LAUNCH_IJK(s, k_flags_from_levelset, f->di(), G(phi)->d);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_levelset_join(mf_grid* a, const mf_grid* b) {
  MF_API_BEGIN
  REQ_GRID(a, G(a)->isReal(), "this", "LevelsetGrid::join");
  REQ_GRID(b, G(b)->isReal(), "o", "LevelsetGrid::join");
  Grid* ga = G(a); Solver* s = ga->s; s->use();
  check_same_solver(ga, G(b), "LevelsetGrid::join");
  MF_LAUNCH_IJK(s, k_ls_join, ga->d, G(b)->d);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_update_fractions(const mf_grid* flags, const mf_grid* phiObs, mf_grid* fractions, int boundaryWidth, float fracThreshold) {
  MF_API_BEGIN
  REQ_GRID(flags, G(flags)->isInt(), "flags", "updateFractions");
  REQ_GRID(phiObs, G(phiObs)->isReal(), "phiObs", "updateFractions");
  REQ_GRID(fractions, G(fractions)->isMAC(), "fractions", "updateFractions");
  const Grid* f = G(flags); Solver* s = f->s; s->use();
  check_same_solver(f, G(phiObs), "updateFractions"); check_same_solver(f, G(fractions), "updateFractions");
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "updateFractions: grid too small");
  MF_LAUNCH_IJK(s, k_update_fractions, f->di(), G(phiObs)->d, v3(G(fractions)), boundaryWidth, fracThreshold);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_set_obstacle_flags(mf_grid* flags, const mf_grid* phiObs, const mf_grid* fractions, const mf_grid* phiOut, const mf_grid* phiIn, int boundaryWidth) {
  MF_API_BEGIN
  REQ_GRID(flags, G(flags)->isInt(), "flags", "setObstacleFlags");
  REQ_GRID(phiObs, G(phiObs)->isReal(), "phiObs", "setObstacleFlags");
  if (fractions) REQ_GRID(fractions, G(fractions)->isMAC(), "fractions", "setObstacleFlags");
  if (phiOut) REQ_GRID(phiOut, G(phiOut)->isReal(), "phiOut", "setObstacleFlags");
  if (phiIn) REQ_GRID(phiIn, G(phiIn)->isReal(), "phiIn", "setObstacleFlags");
  Grid* f = G(flags); Solver* s = f->s; s->use();
  check_same_solver(f, G(phiObs), "setObstacleFlags");
  markFlagsChanged(f);
  CV3 fr; fr.x = fr.y = fr.z = nullptr; if (fractions) fr = cv3(G(fractions));
  MF_LAUNCH_IJK(s, k_set_obstacle_flags, f->di(), fr, fractions ? 1 : 0, G(phiObs)->d, phiOut ? G(phiOut)->d : (const float*)nullptr,
                phiIn ? G(phiIn)->d : (const float*)nullptr, boundaryWidth);
  MF_LAUNCH_CHECK();
  MF_API_END
}

} // extern "C"
