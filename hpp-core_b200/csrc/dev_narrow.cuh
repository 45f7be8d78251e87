// Device narrow phase: closed forms for sphere/capsule/box pairs and GJK on the core
// shapes for everything else.  Replaces coal::ComputeCollision as called through
// hpp::core::CollisionPair::collide (reference include/hpp/core/collision-pair.hh:67-80,
// src/obstacle-user.cc:42-51, src/continuous-validation/body-pair-collision.cc:175-198).
//
// Semantics: d = separation of the two convex shapes (cores
// overlapping => 0 - radii), distance_lower_bound = d - security_margin,
// collision <=> d - security_margin <= 1e-12.  EPA is not needed: hpp-core's validators never
// read a penetration depth.
#pragma once
#include "dev_math.cuh"

namespace hppb200 {

enum { SH_SPHERE = 0, SH_CAPSULE = 1, SH_BOX = 2, SH_CONVEX = 3 };
// pair kinds after canonical ordering (shape type of first <= type of second)
enum { PK_SPHERE_SPHERE = 0, PK_SPHERE_CAPSULE = 1, PK_CAPSULE_CAPSULE = 2, PK_SPHERE_BOX = 3, PK_GJK = 4 };

#define HPPB200_COLLISION_THRESHOLD 1e-12
#define HPPB200_GJK_EPS_REL 1e-10
#define HPPB200_GJK_EPS_ABS2 1e-28
#define HPPB200_GJK_MAX_ITER 64

// 48-byte shape record (lives in the model blob, staged in shared memory)
struct DevShape {
  double p[4];      // sphere: r | capsule: r, halfLength | box: hx hy hz
  int32_t type;
  int32_t vbegin;   // convex: first vertex in the pool
  int32_t vcount;
  int32_t pad;
};

__device__ __forceinline__ double inflation(const DevShape& s) { return s.type <= SH_CAPSULE ? s.p[0] : 0.0; }

// support of the core shape, local frame; ties -> positive side / first maximal vertex
__device__ __forceinline__ V3 supportLocal(const DevShape& s, const double* __restrict__ verts, V3 d) {
  switch (s.type) {
    case SH_SPHERE: return mk(0, 0, 0);
    case SH_CAPSULE: return mk(0, 0, d.z >= 0 ? s.p[1] : -s.p[1]);
    case SH_BOX: return mk(d.x >= 0 ? s.p[0] : -s.p[0], d.y >= 0 ? s.p[1] : -s.p[1], d.z >= 0 ? s.p[2] : -s.p[2]);
    default: {
      int best = 0;
      double bv = -CUDART_INF;
      const double* vp = verts + 3 * (size_t)s.vbegin;
      // four independent dot chains per trip (one job is a dependent FP64 chain; the scan is its longest stretch on convex
      // shapes), compared in vertex order: the first maximal vertex, as before
      int i = 0;
      for (; i + 4 <= s.vcount; i += 4) {
        const double h0 = dot(mk(vp[3 * i], vp[3 * i + 1], vp[3 * i + 2]), d);
        const double h1 = dot(mk(vp[3 * i + 3], vp[3 * i + 4], vp[3 * i + 5]), d);
        const double h2 = dot(mk(vp[3 * i + 6], vp[3 * i + 7], vp[3 * i + 8]), d);
        const double h3 = dot(mk(vp[3 * i + 9], vp[3 * i + 10], vp[3 * i + 11]), d);
        if (h0 > bv) { bv = h0; best = i; }
        if (h1 > bv) { bv = h1; best = i + 1; }
        if (h2 > bv) { bv = h2; best = i + 2; }
        if (h3 > bv) { bv = h3; best = i + 3; }
      }
      for (; i < s.vcount; ++i) {
        double h = dot(mk(vp[3 * i], vp[3 * i + 1], vp[3 * i + 2]), d);
        if (h > bv) { bv = h; best = i; }
      }
      return mk(vp[3 * best], vp[3 * best + 1], vp[3 * best + 2]);
    }
  }
}

// ---- simplex sub-algorithms (closest point to the origin) -------------------------
// Ericson, Real-Time Collision Detection §5.1.5: Voronoi-region tests with p = origin.
// Outputs the supporting sub-simplex (o0..o2, n) and the closest point v.
// Same decisions and the same expressions as the textbook form with its early returns (kept in the oracle,
// oracle_narrowphase.hpp), but every lane evaluates all six region tests and ONE division: in a warp of GJK jobs each
// lane used to leave through a different return (2.8 of 32 lanes active in this function, 18 % of wfNarrowKernel's
// instructions).  The region is the FIRST test that holds, in the original order; outputs a region does not set keep the
// caller's values, as before.
__device__ __forceinline__ void closestTriangle(V3 A, V3 B, V3 C, V3& o0, V3& o1, V3& o2, int& n, V3& v) {
  const V3 ab = B - A, ac = C - A;
  const double d1 = -dot(ab, A), d2 = -dot(ac, A);
  const double d3 = -dot(ab, B), d4 = -dot(ac, B);
  const double vc = fma(d1, d4, -(d3 * d2));
  const double d5 = -dot(ab, C), d6 = -dot(ac, C);
  const double vb = fma(d5, d2, -(d1 * d6));
  const double va = fma(d3, d6, -(d5 * d4));
  const double e1 = d4 - d3, e2 = d5 - d6;
  int reg = 7;
  if (va <= 0 && e1 >= 0 && e2 >= 0) reg = 6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) reg = 5;
  if (d6 >= 0 && d5 <= d6) reg = 4;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) reg = 3;
  if (d3 >= 0 && d4 <= d3) reg = 2;
  if (d1 <= 0 && d2 <= 0) reg = 1;
  const V3 nn = cross(ab, ac);
  double num, den;
  if (reg == 3) { num = d1; den = d1 - d3; }
  else if (reg == 5) { num = d2; den = d2 - d6; }
  else if (reg == 6) { num = e1; den = e1 + e2; }
  else { num = dot(A, nn); den = dot(nn, nn); }
  const double r = num / den;  // (regions 1, 2, 4 do not use it)
  if (reg == 1) { o0 = A; n = 1; v = A; }
  else if (reg == 2) { o0 = B; n = 1; v = B; }
  else if (reg == 4) { o0 = C; n = 1; v = C; }
  else if (reg == 3) { o0 = A; o1 = B; n = 2; v = axpy(r, ab, A); }
  else if (reg == 5) { o0 = A; o1 = C; n = 2; v = axpy(r, ac, A); }
  else if (reg == 6) { o0 = B; o1 = C; n = 2; v = axpy(r, C - B, B); }
  else { o0 = A; o1 = B; o2 = C; n = 3; v = r * nn; }
}

struct GjkCounters {
  unsigned calls, iterations, cap_hits;
};

// GJK as an explicit state + one-iteration step, so that a kernel can either run it to completion
// (gjkCoreDistance) or interleave iterations of many jobs and refill finished lanes (kernels_wavefront.cu).
// The arithmetic is the same expression sequence either way.
struct GjkState {
  V3 w0, w1, w2, v;
  double vv;
  int n, it;
};

__device__ __forceinline__ void gjkBegin(GjkState& s, const T3& rel) {
  s.w0 = mk(0, 0, 0); s.w1 = s.w0; s.w2 = s.w0;
  s.n = 0;
  s.it = 0;
  s.v = -rel.t;
  s.vv = dot(s.v, s.v);
  if (s.vv == 0) { s.v = mk(1, 0, 0); s.vv = 1; }
}

// One iteration.  Returns true when the algorithm has terminated; `out` is then the distance between
// the core shapes (0 when they intersect) and s.it the number of iterations minus one.
// `rel` is a T3 in registers, or a RelStrided view of 12 doubles kept elsewhere (shared memory: 24 registers less
// across the simplex update, where the pressure peaks); the two give the same bits (same dot() order).
template <int STRIDE>
struct RelStrided {
  const double* p;  // entry k (r[0..8], t.x, t.y, t.z) at p[k * STRIDE]
};
__device__ __forceinline__ V3 relRtmul(const T3& R, V3 v) { return rtmul(R, v); }
__device__ __forceinline__ V3 relRmulT(const T3& R, V3 q) { return rmul(R, q) + R.t; }
template <int STRIDE>
__device__ __forceinline__ V3 relRtmul(const RelStrided<STRIDE>& R, V3 v) {
  const double* p = R.p;
  return mk(dot(mk(p[0 * STRIDE], p[3 * STRIDE], p[6 * STRIDE]), v), dot(mk(p[1 * STRIDE], p[4 * STRIDE], p[7 * STRIDE]), v),
            dot(mk(p[2 * STRIDE], p[5 * STRIDE], p[8 * STRIDE]), v));
}
template <int STRIDE>
__device__ __forceinline__ V3 relRmulT(const RelStrided<STRIDE>& R, V3 q) {
  const double* p = R.p;
  return mk(dot(mk(p[0 * STRIDE], p[1 * STRIDE], p[2 * STRIDE]), q), dot(mk(p[3 * STRIDE], p[4 * STRIDE], p[5 * STRIDE]), q),
            dot(mk(p[6 * STRIDE], p[7 * STRIDE], p[8 * STRIDE]), q)) +
         mk(p[9 * STRIDE], p[10 * STRIDE], p[11 * STRIDE]);
}
template <class REL>
__device__ __forceinline__ bool gjkStep(GjkState& s, const DevShape& A, const DevShape& B, const REL& rel,
                                        const double* __restrict__ verts, double& out) {
  const V3 v = s.v;
  const double vv = s.vv;
  const int n = s.n;
  const V3 w0 = s.w0, w1 = s.w1, w2 = s.w2;
  V3 sa = supportLocal(A, verts, -v);
  V3 sbl = supportLocal(B, verts, relRtmul(rel, v));
  V3 sb = relRmulT(rel, sbl);
  V3 w = sa - sb;
  double vw = dot(v, w);
  bool stop = (n > 0 && (vv - vw) <= HPPB200_GJK_EPS_REL * vv);
  bool dup = false;
  if (n > 0 && w0.x == w.x && w0.y == w.y && w0.z == w.z) dup = true;
  if (n > 1 && w1.x == w.x && w1.y == w.y && w1.z == w.z) dup = true;
  if (n > 2 && w2.x == w.x && w2.y == w.y && w2.z == w.z) dup = true;
  if (stop || dup) { out = sqrt(vv); return true; }
  // tentative simplex T = S + {w}
  V3 t0 = w0, t1 = w1, t2 = w2;
  int tn;
  V3 nv;
  bool inside = false;
  if (n == 0) { t0 = w; tn = 1; nv = w; }
  else if (n == 1) {
    // segment A = w0 (old), B = w (new)
    V3 d = w - w0;
    double dd = dot(d, d);
    double sp = -dot(w0, d);
    if (dd == 0 || sp >= dd) { t0 = w; tn = 1; nv = w; }
    else if (sp <= 0) { tn = 1; nv = w0; }
    else { double r = sp / dd; t1 = w; tn = 2; nv = axpy(r, d, w0); }
  } else if (n == 2) {
    closestTriangle(w0, w1, w, t0, t1, t2, tn, nv);
  } else {
    // tetrahedron (w0,w1,w2,w): Ericson §5.1.6, all four faces
    double best = CUDART_INF;
    bool any = false;
    tn = 0; nv = v;
#ifndef HPPB200_GJK_FACE_UNROLL
#define HPPB200_GJK_FACE_UNROLL 1  // 1: one copy of the face code (4: four closestTriangle bodies inline, ~4x the SASS of the tetrahedron case)
#endif
    constexpr int kFaceUnroll = HPPB200_GJK_FACE_UNROLL;
#pragma unroll kFaceUnroll
    for (int f = 0; f < 4; ++f) {
      V3 a, b, c, d;
      if (f == 0) { a = w0; b = w1; c = w2; d = w; }
      else if (f == 1) { a = w0; b = w1; c = w; d = w2; }
      else if (f == 2) { a = w0; b = w2; c = w; d = w1; }
      else { a = w1; b = w2; c = w; d = w0; }
      V3 nn = cross(b - a, c - a);
      double sp = -dot(a, nn);
      double sd = dot(d - a, nn);
      if (sd == 0 || sp * sd < 0) {
        V3 q0, q1, q2, qv;
        q0 = q1 = q2 = a;
        int qn;
        closestTriangle(a, b, c, q0, q1, q2, qn, qv);
        double qq = dot(qv, qv);
        if (qq < best) { best = qq; t0 = q0; t1 = q1; t2 = q2; tn = qn; nv = qv; }
        any = true;
      }
    }
    inside = !any;
  }
  if (inside) { out = 0.0; return true; }
  double nvv = dot(nv, nv);
  if (nvv <= HPPB200_GJK_EPS_ABS2) { out = 0.0; return true; }
  if (n > 0 && nvv >= vv) { out = sqrt(vv); return true; }  // no progress: keep previous closest point
  s.w0 = t0; s.w1 = t1; s.w2 = t2; s.n = tn; s.v = nv; s.vv = nvv;
  s.it += 1;
  if (s.it == HPPB200_GJK_MAX_ITER) { out = sqrt(nvv); return true; }
  return false;
}

// Distance between core shapes; `rel` = placement of B in A's frame.  0 when cores intersect.
static __device__ __noinline__ double gjkCoreDistance(const DevShape& A, const DevShape& B, const T3& rel,
                                               const double* __restrict__ verts, GjkCounters& cnt) {
  GjkState s;
  gjkBegin(s, rel);
  double out;
  while (!gjkStep(s, A, B, rel, verts, out)) {}
  cnt.calls += 1;
  cnt.iterations += s.it + 1;
  if (s.it == HPPB200_GJK_MAX_ITER) cnt.cap_hits += 1;
  return out;
}

__device__ __forceinline__ double clamp01(double x) { return x < 0 ? 0 : (x > 1 ? 1 : x); }

__device__ __forceinline__ double pointSegmentDist2(V3 c, V3 p, V3 u, double h) {
  V3 d = c - p;
  double s = dot(d, u);
  s = s < -h ? -h : (s > h ? h : s);
  V3 e = axpy(-s, u, d);
  return dot(e, e);
}

// Ericson §5.1.9
__device__ __forceinline__ double segmentSegmentDist2(V3 p1, V3 q1, V3 p2, V3 q2) {
  const double EPS = 1e-300;
  V3 d1 = q1 - p1, d2 = q2 - p2, r = p1 - p2;
  double a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r);
  double s, t;
  if (a <= EPS && e <= EPS) return dot(r, r);
  if (a <= EPS) { s = 0; t = clamp01(f / e); }
  else {
    double c = dot(d1, r);
    if (e <= EPS) { t = 0; s = clamp01(-c / a); }
    else {
      // the general case without branches: two divisions on every lane, operands by select (same expressions as the branchy
      // form in the oracle; a warp's lanes used to split three ways at the second one)
      double b = dot(d1, d2);
      double denom = fma(a, e, -(b * b));
      const double s0 = clamp01(fma(b, f, -(c * e)) / denom);
      s = denom > 0 ? s0 : 0;
      double tnom = fma(b, s, f);
      const bool lo = tnom < 0, hi = !lo && tnom > e;
      const double num = lo ? -c : (hi ? b - c : tnom), den = (lo || hi) ? a : e;
      const double q = num / den;
      t = lo ? 0.0 : (hi ? 1.0 : q);
      s = (lo || hi) ? clamp01(q) : s;
    }
  }
  V3 c1 = axpy(s, d1, p1), c2 = axpy(t, d2, p2);
  V3 dd = c1 - c2;
  return dot(dd, dd);
}

// Signed distance (no margin) of a canonically ordered pair.
__device__ __forceinline__ double shapeDistance(int kind, const DevShape& A, const T3& TA, const DevShape& B,
                                                const T3& TB, const double* __restrict__ verts, GjkCounters& cnt) {
  switch (kind) {
    case PK_SPHERE_SPHERE: {
      V3 d = TA.t - TB.t;
      return sqrt(dot(d, d)) - A.p[0] - B.p[0];
    }
    case PK_SPHERE_CAPSULE:
      return sqrt(pointSegmentDist2(TA.t, TB.t, rcol(TB, 2), B.p[1])) - A.p[0] - B.p[0];
    case PK_CAPSULE_CAPSULE: {
      V3 ua = rcol(TA, 2), ub = rcol(TB, 2);
      V3 p1 = axpy(-A.p[1], ua, TA.t), q1 = axpy(A.p[1], ua, TA.t);
      V3 p2 = axpy(-B.p[1], ub, TB.t), q2 = axpy(B.p[1], ub, TB.t);
      return sqrt(segmentSegmentDist2(p1, q1, p2, q2)) - A.p[0] - B.p[0];
    }
    case PK_SPHERE_BOX: {
      V3 c = rtmul(TB, TA.t - TB.t);
      double ex = fabs(c.x) - B.p[0], ey = fabs(c.y) - B.p[1], ez = fabs(c.z) - B.p[2];
      double ox = ex > 0 ? ex : 0, oy = ey > 0 ? ey : 0, oz = ez > 0 ? ez : 0;
      double out2 = fma(oz, oz, fma(oy, oy, ox * ox));
      if (out2 > 0) return sqrt(out2) - A.p[0];
      double m = ex > ey ? ex : ey;
      m = m > ez ? m : ez;
      return m - A.p[0];
    }
    default: {
      T3 rel = composeInv(TA, TB);
      return gjkCoreDistance(A, B, rel, verts, cnt) - inflation(A) - inflation(B);
    }
  }
}

}  // namespace hppb200
