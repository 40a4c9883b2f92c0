// noise.cu -- wavelet noise field and the noise-driven source plugins, SURVEY 8(f) rank 1.
// Reference: WaveletNoiseField (source/noisefield.h:26-97 class, WNoise :160-199, evaluate :302-327; source/noisefield.cpp:
// filter taps :34-38 / :52, downsample :40-48, upsample :54-63, ctor :65-84, generateTile :97-181), MTRand / RandomStream
// (source/util/randomstream.h: randNorm :131-138, getVec3Norm :394), KnApplyNoiseInfl / densityInflow and KnAddNoise / addNoise
// (source/plugin/initplugins.cpp:26-50).
// The 3 x 128^3 tile is produced once per process on the host (a one-off set-up cost, like the reference's static tile) with the
// reference's arithmetic -- MT19937 is std::mt19937, the Gaussian uses the same libm calls -- and uploaded once per solver;
// evaluation is per cell on the device.
#include "common.cuh"
#include "shape_dev.cuh"
#include <cmath>
#include <cstdio>
#include <mutex>
#include <random>

namespace mf {

constexpr int NT = 128;                       // NOISE_TILE_SIZE
constexpr int NT3 = NT * NT * NT;

// ---- host: tile generation ---------------------------------------------------------------------------------------
namespace {

// analysis filter of the wavelet-noise construction (32 taps, symmetric up to the reference's own rounding of two entries)
const float kDown[32] = {
  0.000334f, -0.001528f, 0.000410f, 0.003545f, -0.000938f, -0.008233f, 0.002172f, 0.019120f,
  -0.005040f, -0.044412f, 0.011655f, 0.103311f, -0.025936f, -0.243780f, 0.033979f, 0.655340f,
  0.655340f, 0.033979f, -0.243780f, -0.025936f, 0.103311f, 0.011655f, -0.044412f, -0.005040f,
  0.019120f, 0.002172f, -0.008233f, -0.000938f, 0.003546f, 0.000410f, -0.001528f, 0.000334f };
const float kUp[4] = { 0.25f, 0.75f, 0.75f, 0.25f };

struct Mt {                                   // MTRand's derived distributions on top of the standard generator
  std::mt19937 g;
  explicit Mt(unsigned long seed) : g((std::mt19937::result_type)(seed & 0xffffffffUL)) {}
  double u32() { return (double)g(); }
  double rand() { return u32() * (1.0 / 4294967295.0); }
  double randExc() { return u32() * (1.0 / 4294967296.0); }
  double randDblExc() { return (u32() + 0.5) * (1.0 / 4294967296.0); }
  double randNorm(double mean, double variance) {               // Box-Muller; "variance" scales the radius, as in the reference
    const double r = std::sqrt(-2.0 * std::log(1.0 - randDblExc())) * variance;
    const double phi = 2.0 * 3.14159265358979323846264338328 * randExc();
    return mean + r * std::cos(phi);
  }
};

// periodic analysis step along one line: n -> n/2 coefficients
void lineDown(const float* from, float* to, int n, int stride) {
  for (int i = 0; i < n / 2; i++) {
    float acc = 0.f;
    for (int k = 2 * i - 16; k < 2 * i + 16; k++) acc += kDown[16 + k - 2 * i] * from[(k & (NT - 1)) * stride];
    to[i * stride] = acc;
  }
}
// periodic synthesis step: n/2 -> n; the products are formed in double (0.5 is a double literal in the reference)
void lineUp(const float* from, float* to, int n, int stride) {
  const int h = n / 2;
  for (int i = 0; i < n; i++) {
    float acc = 0.f;
    for (int k = i / 2 - 1; k < i / 2 + 3; k++) {
      int m = k % h; if (m < 0) m += h;
      acc = (float)((double)acc + 0.5 * (double)kUp[k - i / 2 + 1] * (double)from[m * stride]);
    }
    to[i * stride] = acc;
  }
}

void generateTile(float* noise) {
  const int n = NT, n3 = NT3, n3d = 3 * NT3;
  std::vector<float> coarse((size_t)n3d, 0.f), smooth((size_t)n3d, 0.f);
  Mt rs(13322223UL);                                            // WaveletNoiseField::randomSeed
  for (int i = 0; i < n3d; i++) noise[i] = (float)rs.randNorm(0.0, 1.0);
  // band-limit: down- and up-sample along x, then y, then z; `smooth` ends up holding the coarse-scale part
  for (int tile = 0; tile < 3; tile++) {
    const int base = tile * n3;
    for (int iy = 0; iy < n; iy++)
      for (int iz = 0; iz < n; iz++) {
        const int i = base + iy * n + iz * n * n;
        lineDown(&noise[i], &coarse[i], n, 1); lineUp(&coarse[i], &smooth[i], n, 1);
      }
    for (int ix = 0; ix < n; ix++)
      for (int iz = 0; iz < n; iz++) {
        const int i = base + ix + iz * n * n;
        lineDown(&smooth[i], &coarse[i], n, n); lineUp(&coarse[i], &smooth[i], n, n);
      }
    for (int ix = 0; ix < n; ix++)
      for (int iy = 0; iy < n; iy++) {
        const int i = base + ix + iy * n;
        lineDown(&smooth[i], &coarse[i], n, n * n); lineUp(&coarse[i], &smooth[i], n, n * n);
      }
  }
  for (int i = 0; i < n3d; i++) noise[i] -= smooth[i];
  // add an odd-offset copy of the band to itself (evens out the even / odd variance); the copy is written in the reference's
  // counter order, ix slowest
  const int offset = n / 2 + 1;
  size_t cnt = 0;
  for (int tile = 0; tile < 3; tile++)
    for (int ix = 0; ix < n; ix++)
      for (int iy = 0; iy < n; iy++)
        for (int iz = 0; iz < n; iz++)
          coarse[cnt++] = noise[((ix + offset) & (n - 1)) + ((iy + offset) & (n - 1)) * n + ((iz + offset) & (n - 1)) * n * n + tile * n3];
  for (int i = 0; i < n3d; i++) noise[i] += coarse[i];
}

std::mutex gTileLock;
std::vector<float> gTile;                       // process-wide, like WaveletNoiseField::mNoiseTile

const char* const kTileFile = "waveletNoiseTile.bin";

const std::vector<float>& hostTile(bool useFile) {
  std::lock_guard<std::mutex> lk(gTileLock);
  if (!gTile.empty()) return gTile;
  gTile.resize((size_t)3 * NT3);
  if (useFile) {
    if (FILE* fp = fopen(kTileFile, "rb")) {
      const size_t got = fread(gTile.data(), sizeof(float), gTile.size(), fp);
      fclose(fp);
      if (got != gTile.size()) { gTile.clear(); mfErr("Failed to read wavelet noise tile, file invalid/corrupt? (" << kTileFile << ") "); }
      return gTile;
    }
  }
  generateTile(gTile.data());
  if (useFile) {
    if (FILE* fp = fopen(kTileFile, "wb")) { fwrite(gTile.data(), sizeof(float), gTile.size(), fp); fclose(fp); }
  }
  return gTile;
}

} // namespace

struct Noise {
  Solver* s;
  float seed[3];        // mSeedOffset
  float gsInv[3];       // mGsInvX/Y/Z
  float dx;             // FluidSolver::getDx()
};

// ---- device: evaluation ----------------------------------------------------------------------------------------------
struct NoiseD {
  const float* tile;
  float gsInv[3], seed[3], posScale[3], posOffset[3];
  float time;           // getTime() = solver time * dx * timeAnim
  float valOffset, valScale, clampNeg, clampPos;
  int clamp;
};

// quadratic B-spline weights around mid = ceil(p - 1/2)  (noisefield.h:163-181)
__device__ __forceinline__ int splineWeights(float p, float* w) {
  const float q = p - 0.5f;
  const int mid = (int)ceilf(q);
  const float t = (float)mid - q;
  w[0] = t * t * 0.5f;
  w[2] = (1.f - t) * (1.f - t) * 0.5f;
  w[1] = 1.f - w[0] - w[2];
  return mid;
}
__device__ __forceinline__ float wnoise(const float* __restrict__ data, float px, float py, float pz) {
  float wx[3], wy[3], wz[3];
  const int mx = splineWeights(px, wx), my = splineWeights(py, wy), mz = splineWeights(pz, wz);
  float result = 0.f;
#pragma unroll
  for (int z = -1; z <= 1; z++)
#pragma unroll
    for (int y = -1; y <= 1; y++)
#pragma unroll
      for (int x = -1; x <= 1; x++) {
        float weight = 1.0f;
        weight *= wx[x + 1]; weight *= wy[y + 1]; weight *= wz[z + 1];
        const int xc = (mx + x) & (NT - 1), yc = (my + y) & (NT - 1), zc = (mz + z) & (NT - 1);
        result += weight * data[(zc * NT + yc) * NT + xc];
      }
  return result;
}
// WaveletNoiseField::evaluate(pos, tile = 0)
__device__ __forceinline__ float noiseEvaluate(const NoiseD& n, float px, float py, float pz) {
  px *= n.gsInv[0]; py *= n.gsInv[1]; pz *= n.gsInv[2];
  px += n.seed[0]; py += n.seed[1]; pz += n.seed[2];
  px += n.time; py += n.time; pz += n.time;
  px *= n.posScale[0]; py *= n.posScale[1]; pz *= n.posScale[2];
  px += n.posOffset[0]; py += n.posOffset[1]; pz += n.posOffset[2];
  float v = wnoise(n.tile, px, py, pz);
  v += n.valOffset;
  v *= n.valScale;
  if (n.clamp) {
    if (v < n.clampNeg) v = n.clampNeg;
    if (v > n.clampPos) v = n.clampPos;
  }
  return v;
}

#define IJK_ALL \
  const int i = blockIdx.x * blockDim.x + threadIdx.x; \
  const int j = blockIdx.y * blockDim.y + threadIdx.y; \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + g.k0; \
  if (i >= g.nx || j >= g.ny || k >= g.k1) return; \
  const long long idx = (long long)i + (long long)g.nx * j + (long long)g.nx * g.ny * k;

// KnApplyNoiseInfl with the shape's signed distance evaluated in place (densityInflow builds it with computeLevelset)
__global__ void __launch_bounds__(256) k_density_inflow(Geo g, const int* __restrict__ flags, float* __restrict__ density, NoiseD n, ShapeD shape,
                                                        float scale, float sigma) {
  IJK_ALL
  if (!fIsFluid(flags[idx])) return;
  const float sdf = shapeSdf(g, shape, i, j, k);
  if (sdf > sigma) return;
  // clamp(1.0 - 0.5 / sigma * (sdf + sigma), 0.0, 1.0) in double; a NaN (sigma == 0 on the surface) passes through the clamp
  double f = 1.0 - 0.5 / (double)sigma * (double)(sdf + sigma);
  if (f < 0.0) f = 0.0; else if (f > 1.0) f = 1.0;
  const float factor = (float)f;
  const float target = noiseEvaluate(n, (float)i, (float)j, (float)k) * scale * factor;
  if (density[idx] < target) density[idx] = target;
}
// KnAddNoise
__global__ void __launch_bounds__(256) k_add_noise(Geo g, const int* __restrict__ flags, float* __restrict__ density, NoiseD n,
                                                   const float* __restrict__ sdf, float scale) {
  IJK_ALL
  if (!fIsFluid(flags[idx]) || (sdf && sdf[idx] > 0.f)) return;
  density[idx] += noiseEvaluate(n, (float)i, (float)j, (float)k) * scale;
}
__global__ void k_noise_eval_points(NoiseD n, const float* __restrict__ pos, float* __restrict__ out, int count) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < count) out[t] = noiseEvaluate(n, pos[3 * t], pos[3 * t + 1], pos[3 * t + 2]);
}

static const float* deviceTile(Solver* s, bool useFile) {
  if (!s->noiseTile) {
    const std::vector<float>& t = hostTile(useFile);
    CUDA_CHECK(cudaMalloc(&s->noiseTile, sizeof(float) * t.size()));
    CUDA_CHECK(cudaMemcpyAsync(s->noiseTile, t.data(), sizeof(float) * t.size(), cudaMemcpyHostToDevice, s->stream));
    CUDA_CHECK(cudaStreamSynchronize(s->stream));
  }
  return s->noiseTile;
}

static NoiseD noiseD(const Noise* nz, const mf_noise_params* p, float solverTime) {
  mfAssert(nz && p, "NoiseField: null noise object or parameters");
  NoiseD d;
  d.tile = deviceTile(nz->s, false);
  for (int c = 0; c < 3; c++) { d.gsInv[c] = nz->gsInv[c]; d.seed[c] = nz->seed[c]; d.posScale[c] = p->posScale[c]; d.posOffset[c] = p->posOffset[c]; }
  d.time = solverTime * nz->dx * p->timeAnim;          // mParent->getTime() * mParent->getDx() * mTimeAnim
  d.valOffset = p->valOffset; d.valScale = p->valScale; d.clampNeg = p->clampNeg; d.clampPos = p->clampPos; d.clamp = p->clamp;
  return d;
}

} // namespace mf

using namespace mf;

extern "C" {

int mf_noise_generate_tile(float* out) {
  MF_API_BEGIN
  mfAssert(out, "mf_noise_generate_tile: null output");
  generateTile(out);
  MF_API_END
}

mf_noise* mf_noise_create(mf_solver* sp, int fixedSeed, int loadFromFile) {
  try {
    mfAssert(sp, "NoiseField: null solver");
    Solver* s = S(sp); s->use();
    Noise* n = new Noise();
    n->s = s;
    const int mx = std::max(s->nx, std::max(s->ny, s->nz));
    const float scale = (float)(1.0 / (double)mx);
    n->gsInv[0] = scale; n->gsInv[1] = scale; n->gsInv[2] = s->is3D ? scale : 1.f;
    n->dx = scale;
    if (fixedSeed == -1) fixedSeed = 13322223 + 123;
    Mt rs((unsigned long)(long)fixedSeed);
    float a = (float)rs.rand(), b = (float)rs.rand(), c = (float)rs.rand();     // getVec3(): three getReal() in order
    // normalize(Vec3&), vectorbase.h:421-434
    const float l = a * a + b * b + c * c;
    if (std::fabs((double)l - 1.) < (double)(1e-6f * 1e-6f)) { }
    else if (l > 1e-6f * 1e-6f) { const float nrm = std::sqrt(l); const float inv = (float)(1. / (double)nrm); a *= inv; b *= inv; c *= inv; }
    else { a = b = c = 0.f; }
    n->seed[0] = a; n->seed[1] = b; n->seed[2] = c;
    deviceTile(s, loadFromFile != 0);
    return reinterpret_cast<mf_noise*>(n);
  } catch (const std::exception& e) { set_last_error(e.what()); return nullptr; }
}
void mf_noise_destroy(mf_noise* n) { delete reinterpret_cast<Noise*>(n); }

int mf_noise_seed_offset(const mf_noise* n, float* out3) {
  MF_API_BEGIN
  mfAssert(n && out3, "mf_noise_seed_offset: null argument");
  for (int c = 0; c < 3; c++) out3[c] = reinterpret_cast<const Noise*>(n)->seed[c];
  MF_API_END
}

int mf_noise_evaluate(const mf_noise* np, const mf_noise_params* p, float solverTime, const float* hostPos, int count, float* hostOut) {
  MF_API_BEGIN
  const Noise* nz = reinterpret_cast<const Noise*>(np);
  mfAssert(nz && hostPos && hostOut && count >= 0, "mf_noise_evaluate: bad argument");
  Solver* s = nz->s; s->use();
  const NoiseD d = noiseD(nz, p, solverTime);
  if (count == 0) return 0;
  float *dp = nullptr, *dout = nullptr;
  CUDA_CHECK(cudaMalloc(&dp, sizeof(float) * 3 * (size_t)count)); CUDA_CHECK(cudaMalloc(&dout, sizeof(float) * (size_t)count));
  CUDA_CHECK(cudaMemcpyAsync(dp, hostPos, sizeof(float) * 3 * (size_t)count, cudaMemcpyHostToDevice, s->stream));
  MF_LAUNCH(s, k_noise_eval_points, cdiv(count, 256), 256, d, (const float*)dp, dout, count);
  CUDA_CHECK(cudaMemcpyAsync(hostOut, dout, sizeof(float) * (size_t)count, cudaMemcpyDeviceToHost, s->stream));
  CUDA_CHECK(cudaStreamSynchronize(s->stream));
  cudaFree(dp); cudaFree(dout);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_density_inflow(const mf_grid* flags, mf_grid* density, const mf_noise* np, const mf_noise_params* p, const mf_shape* shape,
                      float scale, float sigma, float solverTime) {
  MF_API_BEGIN
  REQ_GRID(flags, G(flags)->isInt(), "flags", "densityInflow"); REQ_GRID(density, G(density)->isReal(), "density", "densityInflow");
  Grid* gr = G(density); Solver* s = gr->s; s->use();
  check_same_solver(G(flags), gr, "densityInflow");
  const Noise* nz = reinterpret_cast<const Noise*>(np);
  mfAssert(nz && nz->s == s, "densityInflow: argument 'noise' is missing or belongs to another solver");
  const NoiseD d = noiseD(nz, p, solverTime);
  const ShapeD sh = shapeOf(shape);
  MF_LAUNCH_IJK(s, k_density_inflow, G(flags)->di(), gr->d, d, sh, scale, sigma);
  MF_LAUNCH_CHECK();
  MF_API_END
}

int mf_add_noise(const mf_grid* flags, mf_grid* density, const mf_noise* np, const mf_noise_params* p, const mf_grid* sdf, float scale,
                 float solverTime) {
  MF_API_BEGIN
  REQ_GRID(flags, G(flags)->isInt(), "flags", "addNoise"); REQ_GRID(density, G(density)->isReal(), "density", "addNoise");
  if (sdf) REQ_GRID(sdf, G(sdf)->isReal(), "sdf", "addNoise");
  Grid* gr = G(density); Solver* s = gr->s; s->use();
  check_same_solver(G(flags), gr, "addNoise");
  const Noise* nz = reinterpret_cast<const Noise*>(np);
  mfAssert(nz && nz->s == s, "addNoise: argument 'noise' is missing or belongs to another solver");
  const NoiseD d = noiseD(nz, p, solverTime);
  MF_LAUNCH_IJK(s, k_add_noise, G(flags)->di(), gr->d, d, sdf ? (const float*)G(sdf)->d : (const float*)nullptr, scale);
  MF_LAUNCH_CHECK();
  MF_API_END
}

} // extern "C"
