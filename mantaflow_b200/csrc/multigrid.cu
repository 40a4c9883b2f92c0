// multigrid.cu -- GridMg on the device: hierarchy construction (setA) and V-cycle (doVCycle).
// Reference: source/multigrid.cpp (ctor :220-319, setA :386-415, doVCycle :448-504, genCoarseGrid :520-578,
// knGenCoarseGridOperator :580-657, smoothGS :713-737, solveCG :796-902).
#include "multigrid.h"
#include "mg_kernels.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>

namespace mf {

struct MgHierarchy {
  Solver* s;
  MgDims D;
  std::vector<LvlD> lv;
  float* b0;                // scaled copy of the rhs, only used when trivial equations exist
  bool aSet;
  bool hasTrivial, nonZeroStencilSum;
  float coarsestAcc, trivialScale;
  int pre, post;
  MgPath* dPaths; int npaths;
  int* dFlags;              // device scratch: [0..1] activate flags, [2..3] coarsening flags
  int usedFallback;
};

// ---- coarsest level: Jacobi-preconditioned CG in double, one CTA (solveCG, multigrid.cpp:796-902) ---------
__device__ __forceinline__ double cgBlockSum(double v, double* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  double t = (l < (int)(blockDim.x >> 5)) ? sh[l] : 0.0;
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  return t;   // every thread holds the block total
}

__device__ __forceinline__ double cgApplyA(const LvlD& L, const MgDims& D, int level0, int v, int vx, int vy, int vz, const double* vec) {
  double sum = 0;
  if (level0) {
    const int n = L.n, Y = L.sx, Z = L.sx * L.sy;
    if (vx > 0)        sum += (double)L.A[n + v - 1] * vec[v - 1];
    if (vx < L.sx - 1) sum += (double)L.A[n + v] * vec[v + 1];
    if (vy > 0)        sum += (double)L.A[2 * n + v - Y] * vec[v - Y];
    if (vy < L.sy - 1) sum += (double)L.A[2 * n + v] * vec[v + Y];
    if (D.is3D) {
      if (vz > 0)        sum += (double)L.A[3 * n + v - Z] * vec[v - Z];
      if (vz < L.sz - 1) sum += (double)L.A[3 * n + v] * vec[v + Z];
    }
    sum += (double)L.A[v] * vec[v];
  } else {
    const int SS = D.SS;
    int s = 0;
    for (int dz = D.zmin; dz <= D.zmax; dz++)
      for (int dy = -1; dy <= 1; dy++)
        for (int dx = -1; dx <= 1; dx++, s++) {
          const int Nx = vx + dx, Ny = vy + dy, Nz = vz + dz;
          if (!mgInGrid(L, Nx, Ny, Nz)) continue;
          const int nn = mgLin(L, Nx, Ny, Nz);
          if (L.type[nn] == vtInactive) continue;
          if (s < SS) sum += (double)L.A[(SS - 1 - s) * L.n + nn] * vec[nn];
          else        sum += (double)L.A[(s - SS + 1) * L.n + v] * vec[nn];
        }
  }
  return sum;
}

__global__ void __launch_bounds__(1024) k_mg_coarse_cg(LvlD L, MgDims D, int level0, float accuracy) {
  __shared__ double sh[32];
  __shared__ double vec[1024];
  const int v = threadIdx.x;
  const bool in = v < L.n;
  const int vx = in ? v % L.sx : 0, vy = in ? (v % (L.sx * L.sy)) / L.sx : 0, vz = in ? v / (L.sx * L.sy) : 0;
  const bool active = in && L.type[v] != vtInactive;
  const double diag = active ? (double)L.A[v] : 1.0;
  double x = in ? (double)L.x[v] : 0.0, r = 0.0, z = 0.0, p = 0.0;
  vec[v] = x;
  __syncthreads();
  if (active) {
    r = (double)L.b[v] - cgApplyA(L, D, level0, v, vx, vy, vz, vec);
    z = r / diag;
    p = z;
  }
  double initialResidual = sqrt(cgBlockSum(active ? r * r : 0.0, sh));
  double alphaTop = cgBlockSum(active ? r * z : 0.0, sh);
  const double acc = (double)accuracy;
  for (int iter = 0; iter < 10000 && initialResidual > 1E-12; iter++) {
    __syncthreads();
    vec[v] = p;
    __syncthreads();
    if (active) z = cgApplyA(L, D, level0, v, vx, vy, vz, vec);
    const double alphaBot = cgBlockSum(active ? p * z : 0.0, sh);
    const double alpha = alphaTop / alphaBot;
    if (active) { x += alpha * p; r -= alpha * z; z = r / diag; }
    const double residual = sqrt(cgBlockSum(active ? r * r : 0.0, sh));
    const double alphaTopNew = cgBlockSum(active ? r * z : 0.0, sh);
    if (residual / initialResidual < acc) break;
    const double beta = alphaTopNew / alphaTop;
    alphaTop = alphaTopNew;
    p = z + beta * p;
  }
  if (in) L.x[v] = (float)x;
}

// ---- host side ----------------------------------------------------------------------------------------------
static void buildPaths(MgHierarchy* mg) {
  // restatement of the path enumeration in GridMg::GridMg (multigrid.cpp:286-318)
  struct P { int N[3], U[3], W[3]; int sc, sf; float rw, iw; bool inU; };
  std::vector<P> paths;
  const int dim = mg->D.dim;
  const int smin[3] = { -1, -1, mg->D.is3D ? -1 : 0 }, smax[3] = { 1, 1, mg->D.is3D ? 1 : 0 };
  const int p7[7][3] = { {0,0,0}, {-1,0,0}, {1,0,0}, {0,-1,0}, {0,1,0}, {0,0,-1}, {0,0,1} };
  const int V[3] = { 1, 1, 1 };
  for (int uz = V[2] * 2 + smin[2]; uz <= V[2] * 2 + smax[2]; uz++)
  for (int uy = V[1] * 2 + smin[1]; uy <= V[1] * 2 + smax[1]; uy++)
  for (int ux = V[0] * 2 + smin[0]; ux <= V[0] * 2 + smax[0]; ux++) {
    const int U[3] = { ux, uy, uz };
    for (int i = 0; i < 1 + 2 * dim; i++) {
      const int W[3] = { U[0] + p7[i][0], U[1] + p7[i][1], U[2] + p7[i][2] };
      for (int nz = W[2] / 2; nz <= (W[2] + 1) / 2; nz++)
      for (int ny = W[1] / 2; ny <= (W[1] + 1) / 2; ny++)
      for (int nx = W[0] / 2; nx <= (W[0] + 1) / 2; nx++) {
        const int sidx = nx + 3 * ny + 9 * nz;
        if (sidx >= 13) {
          P p;
          p.N[0] = nx - 1; p.N[1] = ny - 1; p.N[2] = nz - 1;
          for (int d = 0; d < 3; d++) { p.U[d] = U[d] - V[d] * 2; p.W[d] = W[d] - V[d] * 2; }
          p.sc = sidx - 13; p.sf = (i + 1) / 2; p.inU = (i % 2 == 0);
          p.rw = 1.0f / (float)(1 << ((U[0] % 2) + (U[1] % 2) + (U[2] % 2)));
          p.iw = 1.0f / (float)(1 << ((W[0] % 2) + (W[1] % 2) + (W[2] % 2)));
          paths.push_back(p);
        }
      }
    }
  }
  auto pathLess = [](const P& a, const P& b) {
    const int ka = (a.U[0] + 1) + 3 * (a.U[1] + 1) + 9 * (a.U[2] + 1), kb = (b.U[0] + 1) + 3 * (b.U[1] + 1) + 9 * (b.U[2] + 1);
    if (a.sc == b.sc) return ka < kb;
    return a.sc < b.sc;
  };
  std::sort(paths.begin(), paths.end(), pathLess);
  std::vector<MgPath> hp(paths.size());
  for (size_t q = 0; q < paths.size(); q++) {
    const P& p = paths[q]; MgPath& o = hp[q];
    o.Nx = p.N[0]; o.Ny = p.N[1]; o.Nz = p.N[2]; o.Ux = p.U[0]; o.Uy = p.U[1]; o.Uz = p.U[2];
    o.Wx = p.W[0]; o.Wy = p.W[1]; o.Wz = p.W[2]; o.sc = p.sc; o.sf = p.sf; o.inU = p.inU ? 1 : 0; o.rw = p.rw; o.iw = p.iw;
  }
  mg->npaths = (int)hp.size();
  CUDA_CHECK(cudaMalloc(&mg->dPaths, sizeof(MgPath) * hp.size()));
  CUDA_CHECK(cudaMemcpy(mg->dPaths, hp.data(), sizeof(MgPath) * hp.size(), cudaMemcpyHostToDevice));
}

MgHierarchy* mg_create(Solver* s) {
  mfAssert(s->N < (1LL << 31), "GridMg: grid too large for 32-bit vertex indices");
  MgHierarchy* mg = new MgHierarchy();
  mg->s = s; mg->b0 = nullptr; mg->aSet = false; mg->hasTrivial = false; mg->nonZeroStencilSum = false;
  mg->coarsestAcc = 1E-8f; mg->trivialScale = 1E-6f; mg->pre = 1; mg->post = 1; mg->dPaths = nullptr; mg->npaths = 0; mg->usedFallback = 0;
  const bool is3D = s->nz > 1;
  mg->D.is3D = is3D; mg->D.dim = is3D ? 3 : 2; mg->D.SS = is3D ? 14 : 5; mg->D.SS0 = is3D ? 4 : 3;
  mg->D.zmin = is3D ? -1 : 0; mg->D.zmax = is3D ? 1 : 0;
  int sx = s->nx, sy = s->ny, sz = s->nz;
  long long n = (long long)sx * sy * sz;
  auto addLevel = [&](int l) {
    LvlD L; L.sx = sx; L.sy = sy; L.sz = sz; L.n = (int)n; L.A = L.x = L.b = L.r = nullptr; L.type = nullptr;
    const int ss = (l == 0) ? mg->D.SS0 : mg->D.SS;
    CUDA_CHECK(cudaMalloc(&L.A, sizeof(float) * (size_t)n * ss));
    CUDA_CHECK(cudaMemset(L.A, 0, sizeof(float) * (size_t)n * ss));
    CUDA_CHECK(cudaMalloc(&L.r, sizeof(float) * (size_t)n));
    CUDA_CHECK(cudaMemset(L.r, 0, sizeof(float) * (size_t)n));
    CUDA_CHECK(cudaMalloc(&L.type, (size_t)n));
    if (l > 0) {
      CUDA_CHECK(cudaMalloc(&L.x, sizeof(float) * (size_t)n));
      CUDA_CHECK(cudaMalloc(&L.b, sizeof(float) * (size_t)n));
      CUDA_CHECK(cudaMemset(L.x, 0, sizeof(float) * (size_t)n));
      CUDA_CHECK(cudaMemset(L.b, 0, sizeof(float) * (size_t)n));
    }
    mg->lv.push_back(L);
  };
  addLevel(0);
  for (int l = 1; l <= 100; l++) {          // multigrid.cpp:256-276
    if (sx <= 5 && sy <= 5 && sz <= 5) break;
    if (n <= 1000) break;
    sx = (sx + 2) / 2; sy = (sy + 2) / 2; sz = (sz + 2) / 2;
    n = (long long)sx * sy * sz;
    addLevel(l);
  }
  CUDA_CHECK(cudaMalloc(&mg->dFlags, sizeof(int) * 8));
  buildPaths(mg);
  return mg;
}

void mg_destroy(MgHierarchy* mg) {
  if (!mg) return;
  cudaStreamSynchronize(mg->s->stream);
  for (size_t l = 0; l < mg->lv.size(); l++) {
    LvlD& L = mg->lv[l];
    cudaFree(L.A); cudaFree(L.r); cudaFree(L.type);
    if (l > 0) { cudaFree(L.x); cudaFree(L.b); }
  }
  if (mg->b0) cudaFree(mg->b0);
  cudaFree(mg->dFlags); cudaFree(mg->dPaths);
  delete mg;
}

bool mg_is_a_set(const MgHierarchy* mg) { return mg->aSet; }
void mg_set_coarsest_accuracy(MgHierarchy* mg, float acc) { mg->coarsestAcc = acc; }
void mg_set_smoothing(MgHierarchy* mg, int pre, int post) { mg->pre = pre; mg->post = post; }
int mg_num_levels(const MgHierarchy* mg) { return (int)mg->lv.size(); }
int mg_used_host_fallback(const MgHierarchy* mg) { return mg->usedFallback; }
int mg_level_size(const MgHierarchy* mg, int l, int* sx, int* sy, int* sz) {
  if (l < 0 || l >= (int)mg->lv.size()) return 0;
  *sx = mg->lv[l].sx; *sy = mg->lv[l].sy; *sz = mg->lv[l].sz; return mg->lv[l].n;
}

// serial genCoarseGrid (multigrid.cpp:520-578) on the host: only taken when the order-independent device sweeps leave
// vertices with >= 2 free interpolation vertices.  Bucket min-heap with the reference's LIFO order inside a bucket
// (NKMinHeap, multigrid.cpp:59-186).
namespace {
struct BucketHeap {
  int N, K, size, minKey;
  std::vector<int> key, prev, next;   // K bucket heads followed by N entries
  BucketHeap(int n, int k) : N(n), K(k), size(0), minKey(-1), key(n + k, -1), prev(n + k, -1), next(n + k, -1) {}
  int getKey(int id) const { return key[K + id]; }
  void setKey(int id, int k) {
    const int kid = K + id;
    if (key[kid] == k) return;
    if (key[kid] != -1) {
      const int pred = prev[kid], succ = next[kid];
      next[pred] = succ;
      if (succ != -1) prev[succ] = pred;
      if (key[kid] == minKey) {
        if (size == 1) minKey = -1;
        else for (; minKey < K; minKey++) if (next[minKey] != -1) break;
      }
      size--;
    }
    key[kid] = k;
    if (k == -1) { next[kid] = prev[kid] = -1; return; }
    size++;
    minKey = (minKey == -1) ? k : std::min(minKey, k);
    const int tmp = next[k];
    next[k] = kid; prev[kid] = k; next[kid] = tmp;
    if (tmp != -1) prev[tmp] = kid;
  }
  int popMin() {
    if (size == 0) return -1;
    const int kid = next[minKey], id = kid - K;
    const int pred = prev[kid], succ = next[kid];
    next[pred] = succ;
    if (succ != -1) prev[succ] = pred;
    key[kid] = prev[kid] = next[kid] = -1;
    size--;
    if (size == 0) minKey = -1;
    else for (; minKey < K; minKey++) if (next[minKey] != -1) break;
    return id;
  }
};
}

static void coarsenOnHost(MgHierarchy* mg, int l) {
  Solver* s = mg->s;
  LvlD& F = mg->lv[l - 1]; LvlD& C = mg->lv[l];
  std::vector<unsigned char> ft(F.n), ct(C.n, vtFree);
  CUDA_CHECK(cudaMemcpyAsync(ft.data(), F.type, F.n, cudaMemcpyDeviceToHost, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  BucketHeap heap(F.n, mg->D.is3D ? 9 : 5);
  const int fxy = F.sx * F.sy, cxy = C.sx * C.sy;
  for (int v = 0; v < F.n; v++) {
    if (ft[v] != vtInactive) {
      const int x = v % F.sx, y = (v % fxy) / F.sx, z = v / fxy;
      heap.setKey(v, 1 << ((x % 2) + (y % 2) + (z % 2)));
    }
  }
  while (heap.size > 0) {
    const int v = heap.popMin();
    const int x = v % F.sx, y = (v % fxy) / F.sx, z = v / fxy;
    bool vdone = false;
    for (int iz = z / 2; iz <= (z + 1) / 2; iz++)
    for (int iy = y / 2; iy <= (y + 1) / 2; iy++)
    for (int ix = x / 2; ix <= (x + 1) / 2; ix++) {
      const int i = ix + iy * C.sx + iz * cxy;
      if (ct[i] == vtFree) {
        if (vdone) ct[i] = vtRemoved; else { ct[i] = vtZero; vdone = true; }
        for (int rz = std::max(0, iz * 2 - 1); rz <= std::min(F.sz - 1, iz * 2 + 1); rz++)
        for (int ry = std::max(0, iy * 2 - 1); ry <= std::min(F.sy - 1, iy * 2 + 1); ry++)
        for (int rx = std::max(0, ix * 2 - 1); rx <= std::min(F.sx - 1, ix * 2 + 1); rx++) {
          const int r = rx + ry * F.sx + rz * fxy;
          const int key = heap.getKey(r);
          if (key > 1) heap.setKey(r, key - 1);
          else if (key > -1) heap.setKey(r, -1);
        }
      }
    }
  }
  for (int i = 0; i < C.n; i++) {
    unsigned char c = ct[i];
    if (c == vtFree) c = vtRemoved;
    if (c == vtZero) c = vtActive;
    if (c == vtRemoved) c = vtInactive;
    ct[i] = c;
  }
  CUDA_CHECK(cudaMemcpyAsync(C.type, ct.data(), C.n, cudaMemcpyHostToDevice, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  mg->usedFallback++;
}

static void genCoarseGrid(MgHierarchy* mg, int l) {
  Solver* s = mg->s;
  LvlD& F = mg->lv[l - 1]; LvlD& C = mg->lv[l];
  MF_LAUNCH(s, k_mg_fill_type, cdiv(C.n, 256), 256, C.type, (unsigned char)vtFree, C.n);
  int* flags = mg->dFlags + 2;
  bool pending = false;
  for (int sweep = 0; sweep < 10000; sweep++) {
    CUDA_CHECK(cudaMemsetAsync(flags, 0, sizeof(int) * 2, s->stream));
    MF_LAUNCH(s, k_mg_coarsen_sweep, cdiv(F.n, 256), 256, F, C, flags);
    CUDA_CHECK(cudaMemcpyAsync(s->hostInts, flags, sizeof(int) * 2, cudaMemcpyDeviceToHost, s->stream));
    CUDA_CHECK(cudaStreamSynchronize(s->stream));
    pending = s->hostInts[1] != 0;
    if (!s->hostInts[0]) break;
  }
  if (pending) { coarsenOnHost(mg, l); return; }
  MF_LAUNCH(s, k_mg_coarsen_finalize, cdiv(C.n, 256), 256, C.type, C.n);
}

void mg_set_a(MgHierarchy* mg, const float* A0, const float* Ai, const float* Aj, const float* Ak) {
  Solver* s = mg->s;
  LvlD& L0 = mg->lv[0];
  CUDA_CHECK(cudaMemsetAsync(mg->dFlags, 0, sizeof(int) * 2, s->stream));
  MF_LAUNCH(s, k_mg_copy_activate, cdiv(L0.n, 256), 256, L0, mg->D, A0, Ai, Aj, Ak, mg->trivialScale, mg->dFlags);
  CUDA_CHECK(cudaMemcpyAsync(s->hostInts, mg->dFlags, sizeof(int) * 2, cudaMemcpyDeviceToHost, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  mg->nonZeroStencilSum = s->hostInts[0] != 0;
  mg->hasTrivial = s->hostInts[1] != 0;
  for (size_t l = 1; l < mg->lv.size(); l++) {
    genCoarseGrid(mg, (int)l);
    if (l == 1) { MF_LAUNCH(s, k_mg_galerkin1, cdiv(mg->lv[1].n, 128), 128, mg->lv[0], mg->lv[1], mg->D, mg->dPaths, mg->npaths); }
    else { MF_LAUNCH(s, k_mg_galerkin, cdiv(mg->lv[l].n, 128), 128, mg->lv[l - 1], mg->lv[l], mg->D); }
  }
  MF_LAUNCH_CHECK();
  mg->aSet = true;
}

static void smoothGS(MgHierarchy* mg, int l, bool reversed) {
  Solver* s = mg->s;
  LvlD& L = mg->lv[l];
  if (l == 0) {
    const long long half = (long long)((L.sx + 1) / 2) * L.sy * L.sz;
    for (int c = 0; c < 2; c++) {
      const int color = reversed ? 1 - c : c;
      MF_LAUNCH(s, k_mg_smooth0, cdiv(half, 256), 256, L, mg->D, color);
    }
  } else {
    const int nc = mg->D.is3D ? 8 : 4;
    const int nb = ((L.sx + 1) / 2) * ((L.sy + 1) / 2) * ((L.sz + 1) / 2);
    for (int c = 0; c < nc; c++) {
      const int color = reversed ? nc - 1 - c : c;
      MF_LAUNCH(s, k_mg_smooth, cdiv(nb, 128), 128, L, mg->D, color & 1, (color >> 1) & 1, (color >> 2) & 1);
    }
  }
}

// setRhs + doVCycle(dst) with src == nullptr (multigrid.cpp:426-433, :448-504)
void mg_apply(MgHierarchy* mg, float* dst, const float* rhs) {
  Solver* s = mg->s;
  mfAssert(mg->aSet, "GridMg::setRhs Error: A has not been set.");
  const int maxLevel = (int)mg->lv.size() - 1;
  LvlD& L0 = mg->lv[0];
  if (mg->hasTrivial) {
    if (!mg->b0) CUDA_CHECK(cudaMalloc(&mg->b0, sizeof(float) * (size_t)L0.n));
    MF_LAUNCH(s, k_mg_set_rhs, cdiv(L0.n, 256), 256, L0, rhs, mg->b0, mg->trivialScale);
    L0.b = mg->b0;
  } else L0.b = const_cast<float*>(rhs);
  L0.x = dst;
  CUDA_CHECK(cudaMemsetAsync(dst, 0, sizeof(float) * (size_t)L0.n, s->stream));
  for (int l = 0; l < maxLevel; l++) {
    for (int i = 0; i < mg->pre; i++) smoothGS(mg, l, false);
    LvlD& L = mg->lv[l];
    if (l == 0) { MF_LAUNCH(s, k_mg_residual0, cdiv(L.n, 256), 256, L, mg->D); }
    else { MF_LAUNCH(s, k_mg_residual, cdiv(L.n, 128), 128, L, mg->D); }
    MF_LAUNCH(s, k_mg_restrict, cdiv(mg->lv[l + 1].n, 128), 128, L, mg->lv[l + 1]);
  }
  MF_LAUNCH(s, k_mg_coarse_cg, 1, 1024, mg->lv[maxLevel], mg->D, maxLevel == 0 ? 1 : 0, mg->coarsestAcc);
  for (int l = maxLevel - 1; l >= 0; l--) {
    MF_LAUNCH(s, k_mg_interp_add, cdiv(mg->lv[l].n, 256), 256, mg->lv[l], mg->lv[l + 1]);
    for (int i = 0; i < mg->post; i++) smoothGS(mg, l, true);
  }
  // the trailing calcResidual(0)/calcResidualNorm(0) of doVCycle only feed its return value, which the CG driver
  // ignores (conjugategrad.cpp:162-167) -- skipped.
}

void mg_download_level(MgHierarchy* mg, int l, unsigned char* type, float* A) {
  Solver* s = mg->s;
  LvlD& L = mg->lv[l];
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  if (type) CUDA_CHECK(cudaMemcpy(type, L.type, L.n, cudaMemcpyDeviceToHost));
  if (A) {
    const int ss = (l == 0) ? mg->D.SS0 : mg->D.SS;
    std::vector<float> planes((size_t)L.n * ss);
    CUDA_CHECK(cudaMemcpy(planes.data(), L.A, sizeof(float) * planes.size(), cudaMemcpyDeviceToHost));
    for (int v = 0; v < L.n; v++) for (int e = 0; e < ss; e++) A[(size_t)v * ss + e] = planes[(size_t)e * L.n + v];
  }
}

} // namespace mf
