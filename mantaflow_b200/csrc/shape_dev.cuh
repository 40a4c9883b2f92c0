// shape_dev.cuh -- analytic shapes on the device: inside tests and signed distances of Box, Sphere, Cylinder
// (source/shapes.cpp: Box::isInside :151, BoxSDF :178, Sphere::isInside :240, SphereSDF :309, Cylinder::isInside :324,
// CylinderSDF :367).  Float expressions keep the reference's order.
#pragma once
#include "common.cuh"
#include <cmath>

namespace mf {

struct ShapeD { int type; float a[3]; float b[3]; float r; float zlen; };   // see mf_shape; zlen = |z| of a cylinder (b normalised)

__device__ __forceinline__ float vnorm(float x, float y, float z) {          // norm(Vec3), vectorbase.h:385-389
  const float l = x * x + y * y + z * z;
  if (l <= 1e-6f * 1e-6f) return 0.f;
  return (fabs((double)l - 1.) < (double)(1e-6f * 1e-6f)) ? 1.f : sqrtf(l);
}
__device__ __forceinline__ float fmax2(float a, float b) { return (a < b) ? b : a; }   // std::max
__device__ __forceinline__ float fmin2(float a, float b) { return (b < a) ? b : a; }   // std::min

__device__ __forceinline__ bool shapeInside(const ShapeD& s, float px, float py, float pz) {
  if (s.type == MF_SHAPE_BOX)
    return px >= s.a[0] && py >= s.a[1] && pz >= s.a[2] && px <= s.b[0] && py <= s.b[1] && pz <= s.b[2];
  if (s.type == MF_SHAPE_SPHERE) {
    const float x = (px - s.a[0]) / s.b[0], y = (py - s.a[1]) / s.b[1], z = (pz - s.a[2]) / s.b[2];
    return x * x + y * y + z * z <= s.r * s.r;
  }
  const float dx = px - s.a[0], dy = py - s.a[1], dz = pz - s.a[2];
  const float z = dx * s.b[0] + dy * s.b[1] + dz * s.b[2];
  if (fabsf(z) > s.zlen) return false;
  const float r2 = (dx * dx + dy * dy + dz * dz) - z * z;
  return r2 < s.r * s.r;
}

// BoxSDF / SphereSDF / CylinderSDF
__device__ __forceinline__ float shapeSdf(const Geo& g, const ShapeD& s, int i, int j, int k) {
  const float px = i + 0.5f, py = j + 0.5f, pz = k + 0.5f;
  float out;
  if (s.type == MF_SHAPE_SPHERE) {
    out = vnorm((px - s.a[0]) / s.b[0], (py - s.a[1]) / s.b[1], (pz - s.a[2]) / s.b[2]) - s.r;
  } else if (s.type == MF_SHAPE_CYLINDER) {
    const float dx = px - s.a[0], dy = py - s.a[1], dz = pz - s.a[2];
    const float z = fabsf(dx * s.b[0] + dy * s.b[1] + dz * s.b[2]);
    const float r = sqrtf((dx * dx + dy * dy + dz * dz) - z * z);
    if (z < s.zlen) out = (r < s.r) ? fmax2(r - s.r, z - s.zlen) : r - s.r;
    else if (r < s.r) out = fabsf(z - s.zlen);
    else { const float a = z - s.zlen, b = r - s.r; out = sqrtf(a * a + b * b); }
  } else {
    const float x1 = s.a[0], y1 = s.a[1], z1 = s.a[2], x2 = s.b[0], y2 = s.b[1], z2 = s.b[2];
    const bool inx = px <= x2 && px >= x1, iny = py <= y2 && py >= y1, inz = pz <= z2 && pz >= z1;
    auto sq = [](float a) { return a * a; };
    if (inx && iny && inz) {
      const float mx = fmax2(px - x2, x1 - px), my = fmax2(py - y2, y1 - py);
      float mz = fmax2(pz - z2, z1 - pz);
      if (!g.is3D) mz = mx;
      out = fmax2(mx, fmax2(my, mz));
    } else if (iny && inz) out = fmax2(px - x2, x1 - px);
    else if (inx && inz) out = fmax2(py - y2, y1 - py);
    else if (inx && iny) out = fmax2(pz - z2, z1 - pz);
    else if (px > x1 && px < x2) {
      const float m1 = sqrtf(sq(y1 - py) + sq(z1 - pz)), m2 = sqrtf(sq(y2 - py) + sq(z1 - pz));
      const float m3 = sqrtf(sq(y1 - py) + sq(z2 - pz)), m4 = sqrtf(sq(y2 - py) + sq(z2 - pz));
      out = fmin2(m1, fmin2(m2, fmin2(m3, m4)));
    } else if (py > y1 && py < y2) {
      const float m1 = sqrtf(sq(x1 - px) + sq(z1 - pz)), m2 = sqrtf(sq(x2 - px) + sq(z1 - pz));
      const float m3 = sqrtf(sq(x1 - px) + sq(z2 - pz)), m4 = sqrtf(sq(x2 - px) + sq(z2 - pz));
      out = fmin2(m1, fmin2(m2, fmin2(m3, m4)));
    } else if (pz > x1 && pz < z2) {        // sic: the reference compares p.z with p1.x (shapes.cpp:208)
      const float m1 = sqrtf(sq(y1 - py) + sq(x1 - px)), m2 = sqrtf(sq(y2 - py) + sq(x1 - px));
      const float m3 = sqrtf(sq(y1 - py) + sq(x2 - px)), m4 = sqrtf(sq(y2 - py) + sq(x2 - px));
      out = fmin2(m1, fmin2(m2, fmin2(m3, m4)));
    } else {
      float m = vnorm(px - x1, py - y1, pz - z1);
      m = fmin2(m, vnorm(px - x1, py - y1, pz - z2));
      m = fmin2(m, vnorm(px - x1, py - y2, pz - z1));
      m = fmin2(m, vnorm(px - x1, py - y2, pz - z2));
      m = fmin2(m, vnorm(px - x2, py - y1, pz - z1));
      m = fmin2(m, vnorm(px - x2, py - y1, pz - z2));
      m = fmin2(m, vnorm(px - x2, py - y2, pz - z1));
      m = fmin2(m, vnorm(px - x2, py - y2, pz - z2));
      out = m;
    }
  }
  return out;
}

// host: mf_shape -> ShapeD
inline ShapeD shapeOf(const mf_shape* sh) {
  mfAssert(sh, "Shape: null shape");
  mfAssert(sh->type == MF_SHAPE_BOX || sh->type == MF_SHAPE_SPHERE || sh->type == MF_SHAPE_CYLINDER, "Shape: unknown shape type " << sh->type);
  ShapeD d; d.type = sh->type; d.r = sh->r; d.zlen = 0.f;
  for (int c = 0; c < 3; c++) { d.a[c] = sh->a[c]; d.b[c] = sh->b[c]; }
  if (sh->type == MF_SHAPE_CYLINDER) {
    // Cylinder ctor: mZDir = z; mZ = normalize(mZDir)  (shapes.cpp:316-322, vectorbase.h:421-434)
    float zx = d.b[0], zy = d.b[1], zz = d.b[2];
    const float l = zx * zx + zy * zy + zz * zz;
    if (std::fabs((double)l - 1.) < (double)(1e-6f * 1e-6f)) d.zlen = 1.f;
    else if (l > 1e-6f * 1e-6f) { d.zlen = std::sqrt(l); const float inv = (float)(1. / (double)d.zlen); zx *= inv; zy *= inv; zz *= inv; }
    else { zx = zy = zz = 0.f; d.zlen = 0.f; }
    d.b[0] = zx; d.b[1] = zy; d.b[2] = zz;
  }
  return d;
}


} // namespace mf
