// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing in the product path (hpp-core_b200/,
// include/) may include, link or call this file.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs use it, as the checker / the
// timed CPU baseline.
//
// PARITY UNPINNED for everything in this header: the arithmetic restated here lives in
// third-party libraries that are absent from /root/reference (pinocchio 3.x SE3 algebra,
// glibc sincos).  See SURVEY.md Appendix A.2.
//
// Small dependency-free FP64 vector / rigid-transform algebra + a portable sincos.
// The operation order of every function below is part of the "numerics contract"
// (DESIGN.md §4): the CUDA path evaluates the same expressions in the same order with
// explicit fma() and no implicit contraction, so that results are bit-identical.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace oracle {

struct Vec3 {
  double x, y, z;
};
inline Vec3 operator+(const Vec3& a, const Vec3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(const Vec3& a, const Vec3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator-(const Vec3& a) { return {-a.x, -a.y, -a.z}; }
inline Vec3 operator*(double s, const Vec3& a) { return {s * a.x, s * a.y, s * a.z}; }
// dot(a,b) = fma(az,bz, fma(ay,by, ax*bx))
inline double dot(const Vec3& a, const Vec3& b) {
  return std::fma(a.z, b.z, std::fma(a.y, b.y, a.x * b.x));
}
// cross: each component fma(p,q, -(r*s))
inline Vec3 cross(const Vec3& a, const Vec3& b) {
  return {std::fma(a.y, b.z, -(a.z * b.y)), std::fma(a.z, b.x, -(a.x * b.z)),
          std::fma(a.x, b.y, -(a.y * b.x))};
}
// a + s*d, component-wise fma(s, d, a)
inline Vec3 axpy(double s, const Vec3& d, const Vec3& a) {
  return {std::fma(s, d.x, a.x), std::fma(s, d.y, a.y), std::fma(s, d.z, a.z)};
}

struct Mat3 {
  double m[3][3];
  Vec3 row(int i) const { return {m[i][0], m[i][1], m[i][2]}; }
  Vec3 col(int j) const { return {m[0][j], m[1][j], m[2][j]}; }
};
inline Vec3 mul(const Mat3& R, const Vec3& v) { return {dot(R.row(0), v), dot(R.row(1), v), dot(R.row(2), v)}; }
inline Vec3 tmul(const Mat3& R, const Vec3& v) { return {dot(R.col(0), v), dot(R.col(1), v), dot(R.col(2), v)}; }
inline Mat3 mul(const Mat3& A, const Mat3& B) {
  Mat3 C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i][j] = dot(A.row(i), B.col(j));
  return C;
}
inline Mat3 tmul(const Mat3& A, const Mat3& B) {  // A^T B
  Mat3 C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i][j] = dot(A.col(i), B.col(j));
  return C;
}

// Rigid placement (pinocchio::SE3): x_world = R x_local + t
struct SE3 {
  Mat3 R;
  Vec3 t;
  static SE3 identity() { return {{{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}}, {0, 0, 0}}; }
  static SE3 from12(const double* p) {
    SE3 T;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) T.R.m[i][j] = p[3 * i + j];
    T.t = {p[9], p[10], p[11]};
    return T;
  }
  void to12(double* p) const {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) p[3 * i + j] = R.m[i][j];
    p[9] = t.x; p[10] = t.y; p[11] = t.z;
  }
};
// pinocchio SE3::act : (R1 R2, R1 t2 + t1); translation = dot(row, t2) + t1
inline SE3 compose(const SE3& A, const SE3& B) {
  SE3 C;
  C.R = mul(A.R, B.R);
  Vec3 r = mul(A.R, B.t);
  C.t = {r.x + A.t.x, r.y + A.t.y, r.z + A.t.z};
  return C;
}
// A^{-1} B  (pinocchio SE3::actInv): (Ra^T Rb, Ra^T (tb - ta))
inline SE3 composeInv(const SE3& A, const SE3& B) {
  SE3 C;
  C.R = tmul(A.R, B.R);
  C.t = tmul(A.R, B.t - A.t);
  return C;
}

// ---- portable sin/cos ------------------------------------------------------
// Cody–Waite three-constant reduction by pi/2 followed by the classic minimax kernels
// on [-pi/4, pi/4] (coefficients as published with fdlibm's k_sin.c / k_cos.c,
// Sun Microsystems 1993, freely redistributable).  Valid for |x| < ~1e5, which covers
// joint angles.  Both the oracle and the CUDA path use this instead of libm so that
// forward kinematics is bit-reproducible across host and device.
inline void sincos_portable(double x, double* s, double* c) {
  const double invpio2 = 6.36619772367581382433e-01;
  const double pio2_1 = 1.57079632673412561417e+00;
  const double pio2_2 = 6.07710050630396597660e-11;
  const double pio2_3 = 2.02226624871116645580e-21;
  const double pio2_3t = 8.47842766036889956997e-32;
  double fn = std::nearbyint(x * invpio2);  // round-half-even in the default mode
  // r = x - fn*pio2 in three exact-product steps, tail carried separately
  double r = std::fma(-fn, pio2_1, x);
  r = std::fma(-fn, pio2_2, r);
  double w = fn * pio2_3;
  double y0 = r - w;
  double y1 = std::fma(-fn, pio2_3t, (r - y0) - w);
  int64_t n = (int64_t)fn;
  // kernels on y0 + y1
  double z = y0 * y0;
  // sin kernel
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
               S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
               S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  double v = z * y0;
  double rs = std::fma(z, std::fma(z, std::fma(z, std::fma(z, S6, S5), S4), S3), S2);
  // sin = y0 - ((z*(0.5*y1 - v*rs) - y1) - v*S1)
  double ks = y0 - ((z * std::fma(-v, rs, 0.5 * y1) - y1) - v * S1);
  // cos kernel
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
               C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
               C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  double rc = z * std::fma(z, std::fma(z, std::fma(z, std::fma(z, std::fma(z, C6, C5), C4), C3), C2), C1);
  double hz = 0.5 * z;
  double wc = 1.0 - hz;
  double kc = wc + (((1.0 - wc) - hz) + std::fma(z, rc, -(y0 * y1)));
  switch (n & 3) {
    case 0: *s = ks; *c = kc; break;
    case 1: *s = kc; *c = -ks; break;
    case 2: *s = -ks; *c = -kc; break;
    default: *s = -kc; *c = ks; break;
  }
}

// Portable arctangent (published fdlibm s_atan.c algorithm and coefficients: reduction at
// 7/16, 11/16, 19/16, 39/16 + odd minimax polynomial; Horner steps written with fma).
// Used by the SO(3) logarithm so that Lie-group interpolation is bit-reproducible on host
// and device (only +,-,*,/,fma).  x >= 0 (inf allowed); the callers fold the sign.
inline double atan_portable(double x) {
  const double hi[4] = {4.63647609000806093515e-01, 7.85398163397448278999e-01, 9.82793723247329054082e-01,
                        1.57079632679489655800e+00};
  const double lo[4] = {2.26987774529616870924e-17, 3.06161699786838301793e-17, 1.39033110312309984516e-17,
                        6.12323399573676603587e-17};
  const double a0 = 3.33333333333329318027e-01, a1 = -1.99999999998764832476e-01, a2 = 1.42857142725034663711e-01,
               a3 = -1.11111104054623557880e-01, a4 = 9.09088713343650656196e-02, a5 = -7.69187620504482999495e-02,
               a6 = 6.66107313738753120669e-02, a7 = -5.83357013379057348645e-02, a8 = 4.97687799461593236017e-02,
               a9 = -3.65315727442169155270e-02, a10 = 1.62858201153657823623e-02;
  int id;
  if (x >= 7.378697629483821e19) return hi[3] + lo[3];  // 2^66 and above (and inf)
  if (x < 0.4375) {
    if (x < 7.450580596923828e-09) return x;  // 2^-27
    id = -1;
  } else if (x < 1.1875) {
    if (x < 0.6875) { id = 0; x = (2.0 * x - 1.0) / (2.0 + x); }
    else { id = 1; x = (x - 1.0) / (x + 1.0); }
  } else if (x < 2.4375) { id = 2; x = (x - 1.5) / std::fma(1.5, x, 1.0); }
  else { id = 3; x = -1.0 / x; }
  const double z = x * x, w = z * z;
  const double s1 = z * std::fma(w, std::fma(w, std::fma(w, std::fma(w, std::fma(w, a10, a8), a6), a4), a2), a0);
  const double s2 = w * std::fma(w, std::fma(w, std::fma(w, std::fma(w, a9, a7), a5), a3), a1);
  if (id < 0) return x - x * (s1 + s2);
  return hi[id] - ((x * (s1 + s2) - lo[id]) - x);
}

// ---- unit quaternions (x, y, z, w), Eigen coefficient order ----------------------------
// a * b (Hamilton product)
inline void quatMul(const double* a, const double* b, double* r) {
  const double ax = a[0], ay = a[1], az = a[2], aw = a[3], bx = b[0], by = b[1], bz = b[2], bw = b[3];
  r[0] = std::fma(-az, by, std::fma(ay, bz, std::fma(ax, bw, aw * bx)));
  r[1] = std::fma(-ax, bz, std::fma(az, bx, std::fma(ay, bw, aw * by)));
  r[2] = std::fma(-ay, bx, std::fma(ax, by, std::fma(az, bw, aw * bz)));
  r[3] = std::fma(-az, bz, std::fma(-ay, by, std::fma(-ax, bx, aw * bw)));
}
// SO(3) logarithm of a unit quaternion: theta / |vec| * vec with theta = 2 atan(|vec| / w) folded into
// (-pi, pi] (pinocchio quaternion::log3 as published; series below |vec|^2 = 2^-40).
inline void so3Log(const double* q, double* v) {
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double n2 = std::fma(z, z, std::fma(y, y, x * x));
  double scale;
  if (n2 < 9.094947017729282e-13) {
    const double r2 = n2 / (w * w);
    scale = std::fma(-2.0 / 3.0, r2, 2.0) / w;
  } else {
    const double n = std::sqrt(n2);
    const double half = (w >= 0) ? atan_portable(n / w) : -atan_portable(n / -w);
    scale = (2.0 * half) / n;
  }
  v[0] = scale * x; v[1] = scale * y; v[2] = scale * z;
}
// SO(3) exponential as a unit quaternion (pinocchio quaternion::exp3 as published)
inline void so3Exp(const double* v, double* q) {
  const double t2 = std::fma(v[2], v[2], std::fma(v[1], v[1], v[0] * v[0]));
  double s, c;
  if (t2 < 9.094947017729282e-13) {
    s = std::fma(-1.0 / 48.0, t2, 0.5);
    c = std::fma(-0.125, t2, 1.0);
  } else {
    const double t = std::sqrt(t2);
    double sn;
    sincos_portable(0.5 * t, &sn, &c);
    s = sn / t;
  }
  q[0] = s * v[0]; q[1] = s * v[1]; q[2] = s * v[2]; q[3] = c;
}
// log3(q0^-1 * q1)
inline void so3Difference(const double* q0, const double* q1, double* v) {
  const double c[4] = {-q0[0], -q0[1], -q0[2], q0[3]};
  double r[4];
  quatMul(c, q1, r);
  so3Log(r, v);
}
// Free-flyer in hpp-pinocchio = R^3 x SO(3) (Appendix A.3): q = [x y z, qx qy qz qw], v = [dt, log3].
inline void ffDifference(const double* q0, const double* q1, double* v) {
  v[0] = q1[0] - q0[0]; v[1] = q1[1] - q0[1]; v[2] = q1[2] - q0[2];
  so3Difference(q0 + 3, q1 + 3, v + 3);
}
// q0 (+) u * (q1 (-) q0); the quaternion is renormalised to first order as pinocchio's integrate does
inline void ffInterpolate(const double* q0, const double* q1, double u, double* q) {
  for (int k = 0; k < 3; ++k) q[k] = std::fma(u, q1[k] - q0[k], q0[k]);
  double d[3], e[4], r[4];
  so3Difference(q0 + 3, q1 + 3, d);
  d[0] = u * d[0]; d[1] = u * d[1]; d[2] = u * d[2];
  so3Exp(d, e);
  quatMul(q0 + 3, e, r);
  const double n2 = std::fma(r[3], r[3], std::fma(r[2], r[2], std::fma(r[1], r[1], r[0] * r[0])));
  const double alpha = 0.5 * (3.0 - n2);
  for (int k = 0; k < 4; ++k) q[3 + k] = alpha * r[k];
}
// squaredDistance: translation unweighted, rotation w^2 * |log3|^2
inline double ffSquaredDistance(const double* q0, const double* q1, double w2) {
  const double dx = q1[0] - q0[0], dy = q1[1] - q0[1], dz = q1[2] - q0[2];
  double l[3];
  so3Difference(q0 + 3, q1 + 3, l);
  return std::fma(dz, dz, std::fma(dy, dy, dx * dx)) + w2 * std::fma(l[2], l[2], std::fma(l[1], l[1], l[0] * l[0]));
}

constexpr double kInf = std::numeric_limits<double>::infinity();

}  // namespace oracle
