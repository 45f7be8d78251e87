// Device-side FP64 vector / rigid-transform algebra and the portable sincos.
//
// Numerics contract: every expression below is evaluated in a fixed order
// with explicit fma() and the translation unit is compiled with -fmad=false, so that no
// implicit contraction happens.  The same contract is followed (independently written) by
// the CPU oracle, which is what makes flags and edge parameters bit-comparable.
//
// Replaces on the validation path: pinocchio SE3 algebra used by
// pinocchio::forwardKinematics / updateGeometryPlacements (reference call sites
// src/collision-validation.cc:50-53, src/continuous-validation.cc:134-137).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace hppb200 {

struct V3 {
  double x, y, z;
};
__device__ __forceinline__ V3 mk(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(double s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ double dot(V3 a, V3 b) { return fma(a.z, b.z, fma(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return mk(fma(a.y, b.z, -(a.z * b.y)), fma(a.z, b.x, -(a.x * b.z)), fma(a.x, b.y, -(a.y * b.x)));
}
// a + s*d
__device__ __forceinline__ V3 axpy(double s, V3 d, V3 a) { return mk(fma(s, d.x, a.x), fma(s, d.y, a.y), fma(s, d.z, a.z)); }

// Rigid placement, rotation row-major r[3*i+j]
struct T3 {
  double r[9];
  V3 t;
};
__device__ __forceinline__ V3 rrow(const T3& T, int i) { return mk(T.r[3 * i], T.r[3 * i + 1], T.r[3 * i + 2]); }
__device__ __forceinline__ V3 rcol(const T3& T, int j) { return mk(T.r[j], T.r[3 + j], T.r[6 + j]); }
__device__ __forceinline__ V3 rmul(const T3& T, V3 v) { return mk(dot(rrow(T, 0), v), dot(rrow(T, 1), v), dot(rrow(T, 2), v)); }
__device__ __forceinline__ V3 rtmul(const T3& T, V3 v) { return mk(dot(rcol(T, 0), v), dot(rcol(T, 1), v), dot(rcol(T, 2), v)); }

__device__ __forceinline__ T3 identityT() {
  T3 T;
#pragma unroll
  for (int i = 0; i < 9; ++i) T.r[i] = (i % 4 == 0) ? 1.0 : 0.0;
  T.t = mk(0, 0, 0);
  return T;
}
// A * B : (Ra Rb, Ra tb + ta)
__device__ __forceinline__ T3 compose(const T3& A, const T3& B) {
  T3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.r[3 * i + j] = dot(rrow(A, i), rcol(B, j));
  V3 p = rmul(A, B.t);
  C.t = mk(p.x + A.t.x, p.y + A.t.y, p.z + A.t.z);
  return C;
}
// A^-1 * B : (Ra^T Rb, Ra^T (tb - ta))
__device__ __forceinline__ T3 composeInv(const T3& A, const T3& B) {
  T3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.r[3 * i + j] = dot(rcol(A, i), rcol(B, j));
  C.t = rtmul(A, B.t - A.t);
  return C;
}
__device__ __forceinline__ T3 load12(const double* p) {
  T3 T;
#pragma unroll
  for (int i = 0; i < 9; ++i) T.r[i] = p[i];
  T.t = mk(p[9], p[10], p[11]);
  return T;
}
// strided (SoA) load/store: element k at p[k*stride]
__device__ __forceinline__ T3 load12s(const double* p, size_t stride) {
  T3 T;
#pragma unroll
  for (int i = 0; i < 9; ++i) T.r[i] = p[i * stride];
  T.t = mk(p[9 * stride], p[10 * stride], p[11 * stride]);
  return T;
}
__device__ __forceinline__ void store12s(double* p, size_t stride, const T3& T) {
#pragma unroll
  for (int i = 0; i < 9; ++i) p[i * stride] = T.r[i];
  p[9 * stride] = T.t.x; p[10 * stride] = T.t.y; p[11 * stride] = T.t.z;
}

// Portable sincos: 3-term Cody-Waite reduction by pi/2 + minimax kernels on [-pi/4,pi/4]
// (published fdlibm k_sin/k_cos coefficients).  |x| < ~1e5.  Bit-identical on host and
// device by construction (only +,*,fma,rint).
__device__ __forceinline__ void sincos_portable(double x, double* s, double* c) {
  const double invpio2 = 6.36619772367581382433e-01;
  const double pio2_1 = 1.57079632673412561417e+00;
  const double pio2_2 = 6.07710050630396597660e-11;
  const double pio2_3 = 2.02226624871116645580e-21;
  const double pio2_3t = 8.47842766036889956997e-32;
  double fn = rint(x * invpio2);
  double r = fma(-fn, pio2_1, x);
  r = fma(-fn, pio2_2, r);
  double w = fn * pio2_3;
  double y0 = r - w;
  double y1 = fma(-fn, pio2_3t, (r - y0) - w);
  long long n = (long long)fn;
  double z = y0 * y0;
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
               S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
               S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  double v = z * y0;
  double rs = fma(z, fma(z, fma(z, fma(z, S6, S5), S4), S3), S2);
  double ks = y0 - ((z * fma(-v, rs, 0.5 * y1) - y1) - v * S1);
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
               C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
               C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  double rc = z * fma(z, fma(z, fma(z, fma(z, fma(z, C6, C5), C4), C3), C2), C1);
  double hz = 0.5 * z;
  double wc = 1.0 - hz;
  double kc = wc + (((1.0 - wc) - hz) + fma(z, rc, -(y0 * y1)));
  switch (n & 3) {
    case 0: *s = ks; *c = kc; break;
    case 1: *s = kc; *c = -ks; break;
    case 2: *s = -ks; *c = -kc; break;
    default: *s = -kc; *c = ks; break;
  }
}

// Portable arctangent for x >= 0 (published fdlibm s_atan.c reduction and coefficients, fma Horner steps);
// only +,-,*,/,fma, so host and device agree bit for bit.  Used by the SO(3) logarithm.
__device__ __forceinline__ double atan_portable(double x) {
  if (x >= 7.378697629483821e19) return 1.57079632679489655800e+00 + 6.12323399573676603587e-17;
  double hi = 0, lo = 0;
  bool reduced = true;
  if (x < 0.4375) {
    if (x < 7.450580596923828e-09) return x;
    reduced = false;
  } else if (x < 1.1875) {
    if (x < 0.6875) { hi = 4.63647609000806093515e-01; lo = 2.26987774529616870924e-17; x = (2.0 * x - 1.0) / (2.0 + x); }
    else { hi = 7.85398163397448278999e-01; lo = 3.06161699786838301793e-17; x = (x - 1.0) / (x + 1.0); }
  } else if (x < 2.4375) {
    hi = 9.82793723247329054082e-01; lo = 1.39033110312309984516e-17; x = (x - 1.5) / fma(1.5, x, 1.0);
  } else {
    hi = 1.57079632679489655800e+00; lo = 6.12323399573676603587e-17; x = -1.0 / x;
  }
  const double z = x * x, w = z * z;
  double s1 = fma(w, 1.62858201153657823623e-02, 4.97687799461593236017e-02);
  s1 = fma(w, s1, 6.66107313738753120669e-02);
  s1 = fma(w, s1, 9.09088713343650656196e-02);
  s1 = fma(w, s1, 1.42857142725034663711e-01);
  s1 = z * fma(w, s1, 3.33333333333329318027e-01);
  double s2 = fma(w, -3.65315727442169155270e-02, -5.83357013379057348645e-02);
  s2 = fma(w, s2, -7.69187620504482999495e-02);
  s2 = fma(w, s2, -1.11111104054623557880e-01);
  s2 = w * fma(w, s2, -1.99999999998764832476e-01);
  if (!reduced) return x - x * (s1 + s2);
  return hi - ((x * (s1 + s2) - lo) - x);
}

// ---- R^3 x SO(3) (hpp-pinocchio's free-flyer): q = [x y z, qx qy qz qw], v = [dt, log3] ---------------------
// Replaces hpp-pinocchio LiegroupSpace::{interpolate, operator-} and liegroup squaredDistance for that joint
// (reference call sites src/straight-path.cc:108,138, src/weighed-distance.cc:147-150).
struct Quat {
  double x, y, z, w;
};
__device__ __forceinline__ Quat quatMul(Quat a, Quat b) {
  Quat r;
  r.x = fma(-a.z, b.y, fma(a.y, b.z, fma(a.x, b.w, a.w * b.x)));
  r.y = fma(-a.x, b.z, fma(a.z, b.x, fma(a.y, b.w, a.w * b.y)));
  r.z = fma(-a.y, b.x, fma(a.x, b.y, fma(a.z, b.w, a.w * b.z)));
  r.w = fma(-a.z, b.z, fma(-a.y, b.y, fma(-a.x, b.x, a.w * b.w)));
  return r;
}
// theta / |vec| * vec, theta = 2 atan(|vec| / w) in (-pi, pi]; series below |vec|^2 = 2^-40
__device__ __forceinline__ V3 so3Log(Quat q) {
  const double n2 = fma(q.z, q.z, fma(q.y, q.y, q.x * q.x));
  double scale;
  if (n2 < 9.094947017729282e-13) {
    const double r2 = n2 / (q.w * q.w);
    scale = fma(-2.0 / 3.0, r2, 2.0) / q.w;
  } else {
    const double n = sqrt(n2);
    const double half = (q.w >= 0) ? atan_portable(n / q.w) : -atan_portable(n / -q.w);
    scale = (2.0 * half) / n;
  }
  return mk(scale * q.x, scale * q.y, scale * q.z);
}
__device__ __forceinline__ Quat so3Exp(V3 v) {
  const double t2 = fma(v.z, v.z, fma(v.y, v.y, v.x * v.x));
  double s, c;
  if (t2 < 9.094947017729282e-13) {
    s = fma(-1.0 / 48.0, t2, 0.5);
    c = fma(-0.125, t2, 1.0);
  } else {
    const double t = sqrt(t2);
    double sn;
    sincos_portable(0.5 * t, &sn, &c);
    s = sn / t;
  }
  Quat q;
  q.x = s * v.x; q.y = s * v.y; q.z = s * v.z; q.w = c;
  return q;
}
__device__ __forceinline__ Quat loadQuat(const double* p) { Quat q; q.x = p[0]; q.y = p[1]; q.z = p[2]; q.w = p[3]; return q; }
// log3(q0^-1 q1); q0, q1 point at the 4 quaternion entries.  Out of line: keeps the register budget of the
// kernels that never meet a free-flyer.
static __device__ __noinline__ void so3Difference(const double* q0, const double* q1, double* v) {
  Quat c = loadQuat(q0);
  c.x = -c.x; c.y = -c.y; c.z = -c.z;
  const V3 l = so3Log(quatMul(c, loadQuat(q1)));
  v[0] = l.x; v[1] = l.y; v[2] = l.z;
}
// q1 (-) q0 for one free-flyer (7 entries in, 6 out)
__device__ __forceinline__ void ffDifference(const double* q0, const double* q1, double* v) {
  v[0] = q1[0] - q0[0]; v[1] = q1[1] - q0[1]; v[2] = q1[2] - q0[2];
  so3Difference(q0 + 3, q1 + 3, v + 3);
}
// q0 (+) u (q1 (-) q0): translation by fma, rotation q0 * exp(u log(q0^-1 q1)) renormalised to first order
static __device__ __noinline__ void ffInterpolate(const double* q0, const double* q1, double u, double* q) {
#pragma unroll
  for (int k = 0; k < 3; ++k) q[k] = fma(u, q1[k] - q0[k], q0[k]);
  double d[3];
  so3Difference(q0 + 3, q1 + 3, d);
  const Quat e = so3Exp(mk(u * d[0], u * d[1], u * d[2]));
  const Quat r = quatMul(loadQuat(q0 + 3), e);
  const double n2 = fma(r.w, r.w, fma(r.z, r.z, fma(r.y, r.y, r.x * r.x)));
  const double alpha = 0.5 * (3.0 - n2);
  q[3] = alpha * r.x; q[4] = alpha * r.y; q[5] = alpha * r.z; q[6] = alpha * r.w;
}
// translation unweighted + w^2 |log3|^2
__device__ __forceinline__ double ffSquaredDistance(const double* q0, const double* q1, double w2) {
  const double dx = q1[0] - q0[0], dy = q1[1] - q0[1], dz = q1[2] - q0[2];
  double l[3];
  so3Difference(q0 + 3, q1 + 3, l);
  return fma(dz, dz, fma(dy, dy, dx * dx)) + w2 * fma(l[2], l[2], fma(l[1], l[1], l[0] * l[0]));
}

}  // namespace hppb200
