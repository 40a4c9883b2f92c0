// sample.cuh -- device-side grid sampling with the reference's exact float arithmetic.
// Reference: source/util/interpol.h (BUILD_INDEX :52-69, interpol :72-82, interpolComponent :84-94,
// BUILD_INDEX_SHIFT :116-129, interpolMAC :131-164), source/util/interpolHigh.h (cubicInterp :23-40,
// interpolCubic2D :43-76, interpolCubic :79-172, interpolCubicMAC :174-181),
// source/grid.h (getCentered :460-471, getAtMACX/Y/Z :473-506).
// Association order of every expression follows the reference so results are bit-identical to its
// non-FMA float build (this library is compiled with -fmad=false).
#pragma once
#include "common.cuh"

namespace mf {

struct IdxW {           // result of BUILD_INDEX
  int xi, yi, zi;
  float s0, s1, t0, t1, f0, f1;
};

__device__ __forceinline__ IdxW buildIndex(const Geo& g, float posx, float posy, float posz) {
  IdxW w;
  const float px = posx - 0.5f, py = posy - 0.5f, pz = posz - 0.5f;
  w.xi = (int)px; w.yi = (int)py; w.zi = (int)pz;
  w.s1 = px - (float)w.xi; w.s0 = 1.0f - w.s1;
  w.t1 = py - (float)w.yi; w.t0 = 1.0f - w.t1;
  w.f1 = pz - (float)w.zi; w.f0 = 1.0f - w.f1;
  if (px < 0.0f) { w.xi = 0; w.s0 = 1.0f; w.s1 = 0.0f; }
  if (py < 0.0f) { w.yi = 0; w.t0 = 1.0f; w.t1 = 0.0f; }
  if (pz < 0.0f) { w.zi = 0; w.f0 = 1.0f; w.f1 = 0.0f; }
  if (px >= (float)(g.nx - 1)) { w.xi = g.nx - 2; w.s0 = 0.0f; w.s1 = 1.0f; }
  if (py >= (float)(g.ny - 1)) { w.yi = g.ny - 2; w.t0 = 0.0f; w.t1 = 1.0f; }
  if (g.nz > 1) { if (pz >= (float)(g.nz - 1)) { w.zi = g.nz - 2; w.f0 = 0.0f; w.f1 = 1.0f; } }
  return w;
}

__device__ __forceinline__ float trilerp(const float* __restrict__ d, long long idx, long long Y, long long Z, const IdxW& w) {
  return ((d[idx] * w.t0 + d[idx + Y] * w.t1) * w.s0
        + (d[idx + 1] * w.t0 + d[idx + 1 + Y] * w.t1) * w.s1) * w.f0
       + ((d[idx + Z] * w.t0 + d[idx + Y + Z] * w.t1) * w.s0
        + (d[idx + 1 + Z] * w.t0 + d[idx + 1 + Y + Z] * w.t1) * w.s1) * w.f1;
}

// interpol<Real> / interpolComponent<c> on one SoA plane
__device__ __forceinline__ float interpol(const float* __restrict__ d, const Geo& g, float px, float py, float pz) {
  const IdxW w = buildIndex(g, px, py, pz);
  const long long idx = (long long)w.xi + (long long)g.nx * w.yi + g.Z * w.zi;
  return trilerp(d, idx, g.nx, g.Z, w);
}

// interpolMAC (interpol.h:131-164): returns the three staggered components
__device__ __forceinline__ void interpolMAC(const float* __restrict__ u, const float* __restrict__ v, const float* __restrict__ wv,
                                            const Geo& g, float posx, float posy, float posz, float& rx, float& ry, float& rz) {
  const IdxW w = buildIndex(g, posx, posy, posz);
  int s_xi = (int)posx, s_yi = (int)posy, s_zi = (int)posz;
  float s_s1 = posx - (float)s_xi, s_s0 = 1.0f - s_s1;
  float s_t1 = posy - (float)s_yi, s_t0 = 1.0f - s_t1;
  float s_f1 = posz - (float)s_zi, s_f0 = 1.0f - s_f1;
  if (posx < 0.f) { s_xi = 0; s_s0 = 1.0f; s_s1 = 0.0f; }
  if (posy < 0.f) { s_yi = 0; s_t0 = 1.0f; s_t1 = 0.0f; }
  if (posz < 0.f) { s_zi = 0; s_f0 = 1.0f; s_f1 = 0.0f; }
  if (s_xi >= g.nx - 1) { s_xi = g.nx - 2; s_s0 = 0.0f; s_s1 = 1.0f; }
  if (s_yi >= g.ny - 1) { s_yi = g.ny - 2; s_t0 = 0.0f; s_t1 = 1.0f; }
  if (g.nz > 1) { if (s_zi >= g.nz - 1) { s_zi = g.nz - 2; s_f0 = 0.0f; s_f1 = 1.0f; } }
  const long long X = 1, Y = g.nx, Z = g.Z;
  {
    const float* ref = u + (((long long)w.zi * g.ny + w.yi) * g.nx + s_xi);
    rx = w.f0 * ((ref[0] * w.t0 + ref[Y] * w.t1) * s_s0 + (ref[X] * w.t0 + ref[X + Y] * w.t1) * s_s1)
       + w.f1 * ((ref[Z] * w.t0 + ref[Z + Y] * w.t1) * s_s0 + (ref[X + Z] * w.t0 + ref[X + Y + Z] * w.t1) * s_s1);
  }
  {
    const float* ref = v + (((long long)w.zi * g.ny + s_yi) * g.nx + w.xi);
    ry = w.f0 * ((ref[0] * s_t0 + ref[Y] * s_t1) * w.s0 + (ref[X] * s_t0 + ref[X + Y] * s_t1) * w.s1)
       + w.f1 * ((ref[Z] * s_t0 + ref[Z + Y] * s_t1) * w.s0 + (ref[X + Z] * s_t0 + ref[X + Y + Z] * s_t1) * w.s1);
  }
  {
    const float* ref = wv + (((long long)s_zi * g.ny + w.yi) * g.nx + w.xi);
    rz = s_f0 * ((ref[0] * w.t0 + ref[Y] * w.t1) * w.s0 + (ref[X] * w.t0 + ref[X + Y] * w.t1) * w.s1)
       + s_f1 * ((ref[Z] * w.t0 + ref[Z + Y] * w.t1) * w.s0 + (ref[X + Z] * w.t0 + ref[X + Y + Z] * w.t1) * w.s1);
  }
}

// MAC sampling helpers (grid.h:460-506) -------------------------------------------------
__device__ __forceinline__ void getCentered(const CV3& v, const Geo& g, long long idx, float& cx, float& cy, float& cz) {
  cx = 0.5f * (v.x[idx] + v.x[idx + 1]);
  cy = 0.5f * (v.y[idx] + v.y[idx + g.nx]);
  cz = g.is3D ? 0.5f * (v.z[idx] + v.z[idx + g.Z]) : 0.0f;
}
__device__ __forceinline__ void getAtMACX(const CV3& v, const Geo& g, long long idx, float& cx, float& cy, float& cz) {
  const long long Y = g.nx, Z = g.Z;
  cx = v.x[idx];
  cy = 0.25f * (v.y[idx] + v.y[idx - 1] + v.y[idx + Y] + v.y[idx + Y - 1]);
  cz = g.is3D ? 0.25f * (v.z[idx] + v.z[idx - 1] + v.z[idx + Z] + v.z[idx + Z - 1]) : 0.0f;
}
__device__ __forceinline__ void getAtMACY(const CV3& v, const Geo& g, long long idx, float& cx, float& cy, float& cz) {
  const long long Y = g.nx, Z = g.Z;
  cx = 0.25f * (v.x[idx] + v.x[idx - Y] + v.x[idx + 1] + v.x[idx + 1 - Y]);
  cy = v.y[idx];
  cz = g.is3D ? 0.25f * (v.z[idx] + v.z[idx - Y] + v.z[idx + Z] + v.z[idx + Z - Y]) : 0.0f;
}
// note: getAtMACZ has no 2-D guard in the reference (grid.h:498-506); with Z == 0 the formulas degenerate
__device__ __forceinline__ void getAtMACZ(const CV3& v, const Geo& g, long long idx, float& cx, float& cy, float& cz) {
  const long long Y = g.nx, Z = g.Z;
  cx = 0.25f * (v.x[idx] + v.x[idx - Z] + v.x[idx + 1] + v.x[idx + 1 - Z]);
  cy = 0.25f * (v.y[idx] + v.y[idx - Z] + v.y[idx + Y] + v.y[idx + Y - Z]);
  cz = v.z[idx];
}

// cubic interpolation (interpolHigh.h) -----------------------------------------------------
// Real version: the a2/a3 coefficients are evaluated in double and rounded once (3.0, 2.0 are double
// literals, interpolHigh.h:34-35); the Vec3 version rounds after every scalar*vector product.
template <bool VEC>
__device__ __forceinline__ float cubicInterp(float t, float p0, float p1, float p2, float p3) {
  const float d0 = (p2 - p0) * 0.5f;
  const float d1 = (p3 - p1) * 0.5f;
  const float deltak = p2 - p1;
  const float a0 = p1, a1 = d0;
  float a2, a3;
  if (!VEC) {
    a2 = (float)(3.0 * (double)deltak - 2.0 * (double)d0 - (double)d1);
    a3 = (float)(-2.0 * (double)deltak + (double)d0 + (double)d1);
  } else {
    a2 = ((float)(3.0 * (double)deltak) - (float)(2.0 * (double)d0)) - d1;
    a3 = ((float)(-2.0 * (double)deltak) + d0) + d1;
  }
  const float squared = t * t;
  const float cubed = squared * t;
  return a3 * cubed + a2 * squared + a1 * t + a0;
}

template <bool VEC>
__device__ __forceinline__ float interpolCubic(const float* __restrict__ d, const Geo& g, float posx, float posy, float posz) {
  const float px = posx - 0.5f, py = posy - 0.5f;
  const int x1 = (int)px, y1 = (int)py;
  const int x0 = x1 - 1, x3 = x1 + 2, y0 = y1 - 1, y3 = y1 + 2;
  if (g.Z == 0) {
    if (x0 < 0 || y0 < 0 || x3 >= g.nx || y3 >= g.ny) return interpol(d, g, posx, posy, posz);
    const float xI = px - (float)x1, yI = py - (float)y1;
    float fp[4];
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const float* r = d + (long long)(y0 + b) * g.nx + x0;
      fp[b] = cubicInterp<VEC>(xI, r[0], r[1], r[2], r[3]);
    }
    return cubicInterp<VEC>(yI, fp[0], fp[1], fp[2], fp[3]);
  }
  const float pz = posz - 0.5f;
  const int z1 = (int)pz, z0 = z1 - 1, z3 = z1 + 2;
  if (x0 < 0 || y0 < 0 || z0 < 0 || x3 >= g.nx || y3 >= g.ny || z3 >= g.nz) return interpol(d, g, posx, posy, posz);
  const float xI = px - (float)x1, yI = py - (float)y1, zI = pz - (float)z1;
  float zp[4];
  const long long slab = (long long)g.nx * g.ny;
#pragma unroll
  for (int c = 0; c < 4; c++) {
    float yp[4];
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const float* r = d + (long long)(z0 + c) * slab + (long long)(y0 + b) * g.nx + x0;
      yp[b] = cubicInterp<VEC>(xI, r[0], r[1], r[2], r[3]);
    }
    zp[c] = cubicInterp<VEC>(yI, yp[0], yp[1], yp[2], yp[3]);
  }
  return cubicInterp<VEC>(zI, zp[0], zp[1], zp[2], zp[3]);
}

// normalize(Vec3&) (vectorbase.h:421-434): 1./norm is a double, rounded to Real before the multiply
__device__ __forceinline__ void normalize3(float& x, float& y, float& z) {   // vectorbase.h:421-434
  const float l = x * x + y * y + z * z;
  if (fabs((double)l - 1.) < (double)(1e-6f * 1e-6f)) { /* unit length */ }
  else if (l > 1e-6f * 1e-6f) { const float nrm = sqrtf(l); const float inv = (float)(1. / (double)nrm); x *= inv; y *= inv; z *= inv; }
  else { x = y = z = 0.f; }
}

} // namespace mf
