// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_math.hpp header).
//
// PARITY UNPINNED: restates the *published* narrow-phase algorithms of coal 3.0.x
// (ex hpp-fcl), which is a dependency of the reference that is absent from
// /root/reference.  Call sites in the reference:
//   include/hpp/core/collision-pair.hh:67-80   (CollisionPair::collide)
//   src/obstacle-user.cc:42-51                 (first-hit loop)
//   src/continuous-validation/body-pair-collision.cc:175-198 (distance lower bound)
//
// Semantics restated (Appendix A.4): for a pair of convex shapes
//   d = separation distance (<= 0 when the shapes touch or overlap)
//   distance_lower_bound = d - security_margin
//   isCollision  <=>  d - security_margin <= collision_distance_threshold (1e-12)
// Closed forms: sphere-sphere, sphere-capsule, capsule-capsule (Ericson, Real-Time
// Collision Detection §5.1.9), sphere-box.  Everything else: GJK (Gilbert, Johnson,
// Keerthi 1988) on the *core* shapes (sphere -> point, capsule -> segment) with the
// radii subtracted afterwards, as coal does for swept-sphere shapes.  Penetration depth
// (EPA) is never read by hpp-core's validators, so overlapping cores report d = 0 - radii.
#pragma once
#include "oracle_math.hpp"

namespace oracle {

enum ShapeType { SPHERE = 0, CAPSULE = 1, BOX = 2, CONVEX = 3 };

struct Shape {
  int type = SPHERE;
  double p[4] = {0, 0, 0, 0};      // sphere: r | capsule: r, halfLength | box: hx hy hz
  const double* verts = nullptr;   // convex: vertex pool pointer (xyz triples)
  int nverts = 0;
  double inflation() const { return (type == SPHERE || type == CAPSULE) ? p[0] : 0.0; }
};

constexpr double kCollisionThreshold = 1e-12;  // coal collision_distance_threshold
constexpr double kGjkEpsRel = 1e-10;           // duality-gap tolerance relative to |v|^2
constexpr double kGjkEpsAbs2 = 1e-28;          // |v|^2 below this = cores intersect
constexpr int kGjkMaxIter = 64;

struct GjkStats {
  uint64_t calls = 0, iterations = 0, cap_hits = 0, closed = 0;
};

// Support point of the core shape in its own frame, direction d.
// Ties (component == 0) resolve to the positive side; convex: first maximal vertex.
inline Vec3 supportLocal(const Shape& s, const Vec3& d) {
  switch (s.type) {
    case SPHERE: return {0, 0, 0};
    case CAPSULE: return {0, 0, d.z >= 0 ? s.p[1] : -s.p[1]};
    case BOX: return {d.x >= 0 ? s.p[0] : -s.p[0], d.y >= 0 ? s.p[1] : -s.p[1], d.z >= 0 ? s.p[2] : -s.p[2]};
    default: {
      int best = 0;
      double bv = -kInf;
      for (int i = 0; i < s.nverts; ++i) {
        Vec3 v{s.verts[3 * i], s.verts[3 * i + 1], s.verts[3 * i + 2]};
        double h = dot(v, d);
        if (h > bv) { bv = h; best = i; }
      }
      return {s.verts[3 * best], s.verts[3 * best + 1], s.verts[3 * best + 2]};
    }
  }
}

// ---- closest point to the origin on a simplex ------------------------------
// W holds n points (newest last).  On return W/n hold the smallest sub-simplex that
// supports the closest point v.  Returns true when the origin is inside a tetrahedron.
struct Simplex {
  Vec3 W[4];
  int n = 0;
};

inline void closestSegment(Simplex& S, Vec3& v) {
  const Vec3 A = S.W[0], B = S.W[1];
  Vec3 d = B - A;
  double dd = dot(d, d);
  double s = -dot(A, d);
  if (dd == 0 || s >= dd) { S.W[0] = B; S.n = 1; v = B; return; }
  if (s <= 0) { S.n = 1; v = A; return; }
  double r = s / dd;
  v = axpy(r, d, A);
}

// Ericson §5.1.5 region tests with p = origin; face region projects on the plane normal.
inline void closestTriangle(const Vec3& A, const Vec3& B, const Vec3& C, Simplex& out, Vec3& v) {
  Vec3 ab = B - A, ac = C - A;
  double d1 = -dot(ab, A), d2 = -dot(ac, A);
  if (d1 <= 0 && d2 <= 0) { out.W[0] = A; out.n = 1; v = A; return; }
  double d3 = -dot(ab, B), d4 = -dot(ac, B);
  if (d3 >= 0 && d4 <= d3) { out.W[0] = B; out.n = 1; v = B; return; }
  double vc = std::fma(d1, d4, -(d3 * d2));
  if (vc <= 0 && d1 >= 0 && d3 <= 0) {
    double r = d1 / (d1 - d3);
    out.W[0] = A; out.W[1] = B; out.n = 2; v = axpy(r, ab, A); return;
  }
  double d5 = -dot(ab, C), d6 = -dot(ac, C);
  if (d6 >= 0 && d5 <= d6) { out.W[0] = C; out.n = 1; v = C; return; }
  double vb = std::fma(d5, d2, -(d1 * d6));
  if (vb <= 0 && d2 >= 0 && d6 <= 0) {
    double r = d2 / (d2 - d6);
    out.W[0] = A; out.W[1] = C; out.n = 2; v = axpy(r, ac, A); return;
  }
  double va = std::fma(d3, d6, -(d5 * d4));
  double e1 = d4 - d3, e2 = d5 - d6;
  if (va <= 0 && e1 >= 0 && e2 >= 0) {
    double r = e1 / (e1 + e2);
    out.W[0] = B; out.W[1] = C; out.n = 2; v = axpy(r, C - B, B); return;
  }
  Vec3 n = cross(ab, ac);
  double k = dot(A, n) / dot(n, n);
  out.W[0] = A; out.W[1] = B; out.W[2] = C; out.n = 3;
  v = k * n;
}

// Ericson §5.1.6: test the four faces, keep the nearest; none "outside" => origin inside.
inline bool closestTetrahedron(Simplex& S, Vec3& v) {
  const Vec3 P[4] = {S.W[0], S.W[1], S.W[2], S.W[3]};
  static const int F[4][4] = {{0, 1, 2, 3}, {0, 1, 3, 2}, {0, 2, 3, 1}, {1, 2, 3, 0}};
  bool any = false;
  double best = kInf;
  Simplex bs;
  Vec3 bv{0, 0, 0};
  for (int f = 0; f < 4; ++f) {
    const Vec3 &a = P[F[f][0]], &b = P[F[f][1]], &c = P[F[f][2]], &d = P[F[f][3]];
    Vec3 n = cross(b - a, c - a);
    double sp = -dot(a, n);
    double sd = dot(d - a, n);
    if (sd == 0 || sp * sd < 0) {
      Simplex ts;
      Vec3 tv;
      closestTriangle(a, b, c, ts, tv);
      double q = dot(tv, tv);
      if (q < best) { best = q; bs = ts; bv = tv; }
      any = true;
    }
  }
  if (!any) return true;
  S = bs;
  v = bv;
  return false;
}

// Distance between the core shapes; B's placement relative to A's frame is (R,t).
// Returns 0 when the cores intersect.
inline double gjkCoreDistance(const Shape& A, const Shape& B, const SE3& rel, GjkStats* st) {
  Simplex S;
  Vec3 v = -rel.t;
  double vv = dot(v, v);
  if (vv == 0) { v = {1, 0, 0}; vv = 1; }
  int it = 0;
  for (; it < kGjkMaxIter; ++it) {
    Vec3 dir = -v;
    Vec3 sa = supportLocal(A, dir);
    Vec3 sbl = supportLocal(B, tmul(rel.R, v));
    Vec3 sb = mul(rel.R, sbl) + rel.t;
    Vec3 w = sa - sb;
    double vw = dot(v, w);
    if (S.n > 0 && (vv - vw) <= kGjkEpsRel * vv) break;
    bool dup = false;
    for (int i = 0; i < S.n; ++i)
      if (S.W[i].x == w.x && S.W[i].y == w.y && S.W[i].z == w.z) dup = true;
    if (dup) break;
    Simplex T = S;
    T.W[T.n++] = w;
    Vec3 nv;
    bool inside = false;
    switch (T.n) {
      case 1: nv = w; break;
      case 2: closestSegment(T, nv); break;
      case 3: { Simplex o; closestTriangle(T.W[0], T.W[1], T.W[2], o, nv); T = o; break; }
      default: inside = closestTetrahedron(T, nv); break;
    }
    if (inside) { if (st) { st->calls++; st->iterations += it + 1; } return 0; }
    double nvv = dot(nv, nv);
    if (nvv <= kGjkEpsAbs2) { if (st) { st->calls++; st->iterations += it + 1; } return 0; }
    if (S.n > 0 && nvv >= vv) break;  // no progress: keep the previous closest point
    S = T; v = nv; vv = nvv;
  }
  if (st) { st->calls++; st->iterations += it + 1; if (it == kGjkMaxIter) st->cap_hits++; }
  return std::sqrt(vv);
}

// ---- closed forms -----------------------------------------------------------
inline double clamp01(double x) { return x < 0 ? 0 : (x > 1 ? 1 : x); }

// squared distance point c to segment [p - h*u, p + h*u]
inline double pointSegmentDist2(const Vec3& c, const Vec3& p, const Vec3& u, double h) {
  Vec3 d = c - p;
  double s = dot(d, u);
  s = s < -h ? -h : (s > h ? h : s);
  Vec3 e = axpy(-s, u, d);
  return dot(e, e);
}

// Ericson §5.1.9 ClosestPtSegmentSegment, segments P1Q1 / P2Q2.
inline double segmentSegmentDist2(const Vec3& p1, const Vec3& q1, const Vec3& p2, const Vec3& q2) {
  const double EPS = 1e-300;
  Vec3 d1 = q1 - p1, d2 = q2 - p2, r = p1 - p2;
  double a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r);
  double s, t;
  if (a <= EPS && e <= EPS) { return dot(r, r); }
  if (a <= EPS) { s = 0; t = clamp01(f / e); }
  else {
    double c = dot(d1, r);
    if (e <= EPS) { t = 0; s = clamp01(-c / a); }
    else {
      double b = dot(d1, d2);
      double denom = std::fma(a, e, -(b * b));
      if (denom > 0) s = clamp01(std::fma(b, f, -(c * e)) / denom); else s = 0;
      double tnom = std::fma(b, s, f);
      if (tnom < 0) { t = 0; s = clamp01(-c / a); }
      else if (tnom > e) { t = 1; s = clamp01((b - c) / a); }
      else t = tnom / e;
    }
  }
  Vec3 c1 = axpy(s, d1, p1), c2 = axpy(t, d2, p2);
  Vec3 dd = c1 - c2;
  return dot(dd, dd);
}

// Signed distance of two shapes given world placements (no margin).
inline double shapeDistance(const Shape& A, const SE3& TA, const Shape& B, const SE3& TB, GjkStats* st) {
  if (A.type > B.type) return shapeDistance(B, TB, A, TA, st);
  if (st && !(B.type > CAPSULE && !(A.type == SPHERE && B.type == BOX))) st->closed++;
  if (A.type == SPHERE && B.type == SPHERE) {
    Vec3 d = TA.t - TB.t;
    return std::sqrt(dot(d, d)) - A.p[0] - B.p[0];
  }
  if (A.type == SPHERE && B.type == CAPSULE) {
    return std::sqrt(pointSegmentDist2(TA.t, TB.t, TB.R.col(2), B.p[1])) - A.p[0] - B.p[0];
  }
  if (A.type == CAPSULE && B.type == CAPSULE) {
    Vec3 ua = TA.R.col(2), ub = TB.R.col(2);
    Vec3 p1 = axpy(-A.p[1], ua, TA.t), q1 = axpy(A.p[1], ua, TA.t);
    Vec3 p2 = axpy(-B.p[1], ub, TB.t), q2 = axpy(B.p[1], ub, TB.t);
    return std::sqrt(segmentSegmentDist2(p1, q1, p2, q2)) - A.p[0] - B.p[0];
  }
  if (A.type == SPHERE && B.type == BOX) {
    Vec3 c = tmul(TB.R, TA.t - TB.t);  // centre in the box frame
    double ex = std::fabs(c.x) - B.p[0], ey = std::fabs(c.y) - B.p[1], ez = std::fabs(c.z) - B.p[2];
    double ox = ex > 0 ? ex : 0, oy = ey > 0 ? ey : 0, oz = ez > 0 ? ez : 0;
    double out2 = std::fma(oz, oz, std::fma(oy, oy, ox * ox));
    if (out2 > 0) return std::sqrt(out2) - A.p[0];
    double m = ex > ey ? ex : ey;
    m = m > ez ? m : ez;  // <= 0: centre inside the box, depth to nearest face
    return m - A.p[0];
  }
  SE3 rel = composeInv(TA, TB);
  return gjkCoreDistance(A, B, rel, st) - A.inflation() - B.inflation();
}

}  // namespace oracle
