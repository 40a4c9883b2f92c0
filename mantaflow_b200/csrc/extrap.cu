// extrap.cu -- the simple (sweep-based) extrapolation plugins, SURVEY 8(f) rank 2.
// Reference: source/fastmarch.cpp (knExtrapolateMACSimple :231, knExtrapolateIntoBnd :260, getNormal :304,
// knUnprojectNormalComp :321, extrapolateMACSimple :337, knExtrapolateMACFromWeight :380, extrapolateMACFromWeight :410,
// knExtrapolateLsSimple :439, knSetRemaining :462, extrapolateLsSimple :471, extrapolateVec3Simple :525).
// Every sweep d reads only cells marked d and writes only cells marked 0 (-> d+1), so the in-place sweeps of the reference are
// order-independent and map to one launch per (component, distance); the neighbour sum keeps the reference's order
// +x,-x,+y,-y,+z,-z.  Marks live in a pooled Real grid reinterpreted as int (the reference's Grid<int> tmp).
#include "common.cuh"
#include "sample.cuh"

namespace mf {

#define IJK_BND1 \
  const int i = blockIdx.x * blockDim.x + threadIdx.x; \
  const int j = blockIdx.y * blockDim.y + threadIdx.y; \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0; \
  if (i >= g.nx || j >= g.ny || k >= g.k1) return; \
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k; \
  const bool interior = i >= 1 && j >= 1 && i < g.nx - 1 && j < g.ny - 1 && (!g.is3D || (k >= 1 && k < g.nz - 1));

// tmp.clear() + the marking loop of extrapolateMACSimple (:348-361) for component c (offset = stride of that axis)
__global__ void __launch_bounds__(256) k_ex_mark_mac(Geo g, const int* __restrict__ flags, int* __restrict__ tmp, long long dirOff, int intoObs) {
  IJK_BND1
  int m = 0;
  if (interior) {
    const int f = flags[idx], fm = flags[idx - dirOff];
    if (!intoObs) m = (fIsFluid(f) || fIsFluid(fm)) ? 1 : 0;
    else m = ((fIsFluid(f) || fIsFluid(fm)) && !fIsObstacle(f) && !fIsObstacle(fm)) ? 1 : 0;
  }
  tmp[idx] = m;
}
// average of the neighbours marked d; returns the count (reference order of the six neighbours)
template <class T>
__device__ __forceinline__ int exNeighbourSum(const Geo& g, const T* __restrict__ mark, T d, const float* __restrict__ val, long long idx, float& sum) {
  const long long Y = g.nx, Z = g.Z;
  int nbs = 0; sum = 0.f;
  if (mark[idx + 1] == d) { sum += val[idx + 1]; nbs++; }
  if (mark[idx - 1] == d) { sum += val[idx - 1]; nbs++; }
  if (mark[idx + Y] == d) { sum += val[idx + Y]; nbs++; }
  if (mark[idx - Y] == d) { sum += val[idx - Y]; nbs++; }
  if (g.is3D) {
    if (mark[idx + Z] == d) { sum += val[idx + Z]; nbs++; }
    if (mark[idx - Z] == d) { sum += val[idx - Z]; nbs++; }
  }
  return nbs;
}
// knExtrapolateMACSimple (:231-258) on one SoA component
__global__ void __launch_bounds__(256) k_ex_sweep(Geo g, float* val, int* tmp, int d) {
  IJK_BND1
  if (!interior || tmp[idx] != 0) return;
  float sum;
  const int nbs = exNeighbourSum<int>(g, tmp, d, val, idx, sum);
  if (nbs > 0) { tmp[idx] = d + 1; val[idx] = sum / (float)nbs; }
}
// knExtrapolateMACFromWeight (:380-402): the weight component itself holds the marks (floats 0, 1, 2, ...)
__global__ void __launch_bounds__(256) k_ex_weight_reset(Geo g, float* w) {
  IJK_BND1
  if (interior && w[idx] > 0.f) w[idx] = 1.0f;
}
__global__ void __launch_bounds__(256) k_ex_sweep_weight(Geo g, float* val, float* w, int d) {
  IJK_BND1
  if (!interior || w[idx] != 0.f) return;
  float sum;
  const int nbs = exNeighbourSum<float>(g, w, (float)d, val, idx, sum);
  if (nbs > 0) { w[idx] = (float)(d + 1); val[idx] = sum / (float)nbs; }
}
// knExtrapolateLsSimple<S> (:439-459): several components share the marks; `direction` is added (0 for Vec3)
template <int NC>
__global__ void __launch_bounds__(256) k_ex_sweep_ls(Geo g, float* val, long long np, int* tmp, int d, float direction) {
  IJK_BND1
  if (!interior || tmp[idx] != 0) return;
  float sum[NC]; int nbs = 0;
#pragma unroll
  for (int c = 0; c < NC; c++) nbs = exNeighbourSum<int>(g, tmp, d, val + c * np, idx, sum[c]);
  if (nbs > 0) {
    tmp[idx] = d + 1;
#pragma unroll
    for (int c = 0; c < NC; c++) val[idx + c * np] = sum[c] / (float)nbs + direction;
  }
}
// marks of extrapolateLsSimple / extrapolateVec3Simple: 1 where phi is inside (outside), bnd 1 (0 with include_walls)
__global__ void __launch_bounds__(256) k_ex_mark_phi(Geo g, const float* __restrict__ phi, int* __restrict__ tmp, int inside, int bnd0) {
  IJK_BND1
  int m = 0;
  if (interior || bnd0) { const float p = phi[idx]; m = inside ? (p > 0.f) : (p < 0.f); }
  tmp[idx] = m;
}
// "+ first layer around" (:499-508): unmarked interior cells with a neighbour marked 1 become 2
__global__ void __launch_bounds__(256) k_ex_first_layer(Geo g, int* tmp) {
  IJK_BND1
  if (!interior || tmp[idx] != 0) return;
  const long long Y = g.nx, Z = g.Z;
  bool any = tmp[idx + 1] == 1 || tmp[idx - 1] == 1 || tmp[idx + Y] == 1 || tmp[idx - Y] == 1;
  if (g.is3D) any = any || tmp[idx + Z] == 1 || tmp[idx - Z] == 1;
  if (any) tmp[idx] = 2;
}
// knSetRemaining<S> (:462-467)
template <int NC>
__global__ void __launch_bounds__(256) k_ex_set_remaining(Geo g, float* val, long long np, const int* __restrict__ tmp, float v) {
  IJK_BND1
  if (!interior || tmp[idx] != 0) return;
#pragma unroll
  for (int c = 0; c < NC; c++) val[idx + c * np] = v;
}
// knUnprojectNormalComp (:321-333) with getNormal (:304-320)
__global__ void __launch_bounds__(256) k_ex_unproject(Geo g, V3 vel, const float* __restrict__ phi, float maxDist) {
  IJK_BND1
  if (!interior) return;
  const float p = phi[idx];
  if (p > 0.f || p < -maxDist) return;
  int ci = i, cj = j, ck = k;
  if (ci > g.nx - 2) ci = g.nx - 2; if (ci < 1) ci = 1;
  if (cj > g.ny - 2) cj = g.ny - 2; if (cj < 1) cj = 1;
  long long kd = g.Z;
  if (g.is3D) { if (ck > g.nz - 2) ck = g.nz - 2; if (ck < 1) ck = 1; } else kd = 0;
  const long long c = (long long)ci + (long long)g.nx * cj + (long long)g.nx * g.ny * ck;
  float nx = phi[c + 1] - phi[c - 1], ny = phi[c + g.nx] - phi[c - g.nx], nz = phi[c + kd] - phi[c - kd];
  const float vx = vel.x[idx], vy = vel.y[idx], vz = vel.z[idx];
  if (nx * vx + ny * vy + nz * vz < 0.f) {
    normalize3(nx, ny, nz);
    const float l = nx * vx + ny * vy + nz * vz;
    vel.x[idx] = vx - nx * l; vel.y[idx] = vy - ny * l; vel.z[idx] = vz - nz * l;
  }
}
// knExtrapolateIntoBnd (:260-301): reads the copy velTmp, writes vel on the domain sides
__global__ void __launch_bounds__(256) k_ex_into_bnd(Geo g, const int* __restrict__ flags, V3 vel, CV3 t) {
  IJK_BND1
  (void)interior;
  int c = 0;
  float vx = 0.f, vy = 0.f, vz = 0.f;
  const bool isObs = fIsObstacle(flags[idx]);
  const long long Y = g.nx, Z = g.Z;
  if (i == 0) { vx = t.x[idx + 1]; vy = t.y[idx + 1]; vz = t.z[idx + 1]; if (isObs && vx < 0.f) vx = 0.f; c++; }
  else if (i == g.nx - 1) { vx = t.x[idx - 1]; vy = t.y[idx - 1]; vz = t.z[idx - 1]; if (isObs && vx > 0.f) vx = 0.f; c++; }
  if (j == 0) { vx = t.x[idx + Y]; vy = t.y[idx + Y]; vz = t.z[idx + Y]; if (isObs && vy < 0.f) vy = 0.f; c++; }
  else if (j == g.ny - 1) { vx = t.x[idx - Y]; vy = t.y[idx - Y]; vz = t.z[idx - Y]; if (isObs && vy > 0.f) vy = 0.f; c++; }
  if (g.is3D) {
    if (k == 0) { vx = t.x[idx + Z]; vy = t.y[idx + Z]; vz = t.z[idx + Z]; if (isObs && vz < 0.f) vz = 0.f; c++; }
    else if (k == g.nz - 1) { vx = t.x[idx - Z]; vy = t.y[idx - Z]; vz = t.z[idx - Z]; if (isObs && vz > 0.f) vz = 0.f; c++; }
  }
  if (c > 0) { const float fc = (float)c; vel.x[idx] = vx / fc; vel.y[idx] = vy / fc; vel.z[idx] = vz / fc; }
}

static void copyVecSlabs(Solver* s, float* dst, const float* src) {
  MF_RANKS(s, r) {
    const Span sp = span(s, r);
    if (!sp.n) continue;
    for (int c = 0; c < 3; c++)
      CUDA_CHECK(cudaMemcpyAsync(dst + c * s->Np + sp.off, src + c * s->Np + sp.off, sizeof(float) * (size_t)sp.n, cudaMemcpyDeviceToDevice, s->stream));
  }
  s->rankSync();
}

} // namespace mf

using namespace mf;

extern "C" {

int mf_extrapolate_mac_simple(const mf_grid* flagsp, mf_grid* velp, int distance, const mf_grid* phiObs, int intoObs) {
  MF_API_BEGIN
  REQ_GRID(flagsp, G(flagsp)->isInt(), "flags", "extrapolateMACSimple");
  REQ_GRID(velp, G(velp)->isMAC(), "vel", "extrapolateMACSimple");
  if (phiObs) REQ_GRID(phiObs, G(phiObs)->isReal(), "phiObs", "extrapolateMACSimple");
  const Grid* flags = G(flagsp); Grid* vel = G(velp); Solver* s = flags->s; s->use();
  check_same_solver(flags, vel, "extrapolateMACSimple");
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "extrapolateMACSimple: grid too small");
  TmpReal tmp(s);
  int* t = reinterpret_cast<int*>(tmp.p);
  const int dim = s->is3D ? 3 : 2;
  const long long strides[3] = { 1, (long long)s->nx, (long long)s->nx * s->ny };
  for (int c = 0; c < dim; c++) {
    MF_LAUNCH_IJK(s, k_ex_mark_mac, flags->di(), t, strides[c], intoObs ? 1 : 0);
    for (int d = 1; d < 1 + distance; d++) MF_LAUNCH_IJK(s, k_ex_sweep, vel->comp(c), t, d);
  }
  if (phiObs) MF_LAUNCH_IJK(s, k_ex_unproject, v3(vel), G(phiObs)->d, (float)distance);
  TmpVec velTmp(s);
  copyVecSlabs(s, velTmp.p, vel->d);
  CV3 ct; ct.x = velTmp.comp(0); ct.y = velTmp.comp(1); ct.z = velTmp.comp(2);
  MF_LAUNCH_IJK(s, k_ex_into_bnd, flags->di(), v3(vel), ct);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_extrapolate_mac_from_weight(mf_grid* velp, mf_grid* weightp, int distance) {
  MF_API_BEGIN
  REQ_GRID(velp, G(velp)->isMAC(), "vel", "extrapolateMACFromWeight");
  REQ_GRID(weightp, G(weightp)->isVec(), "weight", "extrapolateMACFromWeight");
  Grid* vel = G(velp); Grid* w = G(weightp); Solver* s = vel->s; s->use();
  check_same_solver(vel, w, "extrapolateMACFromWeight");
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "extrapolateMACFromWeight: grid too small");
  const int dim = s->is3D ? 3 : 2;
  for (int c = 0; c < dim; c++) {
    MF_LAUNCH_IJK(s, k_ex_weight_reset, w->comp(c));
    for (int d = 1; d < 1 + distance; d++) MF_LAUNCH_IJK(s, k_ex_sweep_weight, vel->comp(c), w->comp(c), d);
  }
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_extrapolate_ls_simple(mf_grid* phip, int distance, int inside, int includeWalls) {
  MF_API_BEGIN
  REQ_GRID(phip, G(phip)->isReal(), "phi", "extrapolateLsSimple");
  Grid* phi = G(phip); Solver* s = phi->s; s->use();
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "extrapolateLsSimple: grid too small");
  TmpReal tmp(s);
  int* t = reinterpret_cast<int*>(tmp.p);
  const float direction = inside ? -1.f : 1.f;
  MF_LAUNCH_IJK(s, k_ex_mark_phi, phi->d, t, inside ? 1 : 0, (inside && includeWalls) ? 1 : 0);
  MF_LAUNCH_IJK(s, k_ex_first_layer, t);
  for (int d = 2; d < 1 + distance; d++) MF_LAUNCH_IJK(s, k_ex_sweep_ls<1>, phi->d, 0LL, t, d, direction);
  MF_LAUNCH_IJK(s, k_ex_set_remaining<1>, phi->d, 0LL, t, direction * (float)(distance + 2));
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_extrapolate_vec3_simple(mf_grid* velp, const mf_grid* phip, int distance, int inside) {
  MF_API_BEGIN
  REQ_GRID(velp, G(velp)->isVec(), "vel", "extrapolateVec3Simple");
  REQ_GRID(phip, G(phip)->isReal(), "phi", "extrapolateVec3Simple");
  Grid* vel = G(velp); const Grid* phi = G(phip); Solver* s = vel->s; s->use();
  check_same_solver(vel, phi, "extrapolateVec3Simple");
  mfAssert(s->nx >= 3 && s->ny >= 3 && (!s->is3D || s->nz >= 3), "extrapolateVec3Simple: grid too small");
  TmpReal tmp(s);
  int* t = reinterpret_cast<int*>(tmp.p);
  MF_LAUNCH_IJK(s, k_ex_mark_phi, phi->d, t, inside ? 1 : 0, 0);
  MF_LAUNCH_IJK(s, k_ex_first_layer, t);
  for (int d = 2; d < 1 + distance; d++) MF_LAUNCH_IJK(s, k_ex_sweep_ls<3>, vel->d, s->Np, t, d, 0.f);
  MF_LAUNCH_IJK(s, k_ex_set_remaining<3>, vel->d, s->Np, t, 0.f);
  MF_LAUNCH_CHECK();
  MF_API_END
}

} // extern "C"
