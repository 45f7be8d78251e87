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

}  // namespace hppb200
