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

constexpr double kInf = std::numeric_limits<double>::infinity();

}  // namespace oracle
