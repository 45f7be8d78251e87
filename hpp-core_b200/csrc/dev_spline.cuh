// path::Spline<BernsteinBasis, 3> on the device (reference src/path/spline.cc, include/hpp/core/path/spline.hh): the
// path type of the reference's non-straight continuous-validation test (tests/test-continuous-validation.cc:288-427).
//   q(t) = base (+) P^T b(u),  u = (t - t0) / (t1 - t0) clamped to [0, 1],  b = the four cubic Bernstein polynomials,
//   P [4][nv] tangent vectors at `base`.
// Vector-space configuration spaces only (nq == nv, configuration index == velocity index: the (+) is a sum).
// Operation order = oracle/oracle_validation.hpp SplinePath (same bits).
#pragma once
#include "dev_model.cuh"

namespace hppb200 {

// spline_basis_function<BernsteinBasis, 3>::derivative (src/path/spline.cc:131-159): order-k derivative of the basis at t
__device__ __forceinline__ void splineBasisDerivative(int k, double t, double (&res)[4]) {
  constexpr int Degree = 3;
  const double factors[4] = {1, 1, 2, 6};
  const int binom[4][4] = {{1, 0, 0, 0}, {1, 1, 0, 0}, {1, 2, 1, 0}, {1, 3, 3, 1}};
#pragma unroll
  for (int i = 0; i < 4; ++i) res[i] = 0;
  if (k > Degree) return;
  double pT[4], pO[4];
  pT[0] = 1; pO[0] = 1;
  const double oneMinusT = 1 - t;
#pragma unroll
  for (int i = 1; i < 4; ++i) { pT[i] = pT[i - 1] * t; pO[i] = pO[i - 1] * oneMinusT; }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p0 = (k + i - Degree) > 0 ? (k + i - Degree) : 0, p1 = i < k ? i : k;
    for (int p = p0; p <= p1; ++p) {
      const int ip = i - p;
      res[i] += (double)(((k - p) % 2 == 0 ? 1 : -1) * binom[k][p]) * pT[ip] * pO[Degree - k - ip] / (double)(factors[ip] * factors[Degree - k - ip]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) res[i] *= factors[Degree];
}

// spline_basis_function<BernsteinBasis, 3>::bound (src/path/spline.cc:204-247) at Degree = 3 (b_up(1) = b_up(2) = 1)
__device__ __forceinline__ void splineBasisBound(double u0, double u1, double (&ub)[4]) {
  constexpr int Degree = 3;
  double b0[4], b1[4];
  splineBasisDerivative(1, u0, b0);
  splineBasisDerivative(1, u1, b1);
#pragma unroll
  for (int i = 0; i < 4; ++i) ub[i] = fmax(fabs(b0[i]), fabs(b1[i]));
  if (u0 * Degree < 2 && 2 < u1 * Degree) ub[1] = fmax(ub[1], 1.0);
  if (u0 * Degree < Degree - 2 && Degree - 2 < u1 * Degree) ub[Degree - 1] = fmax(ub[Degree - 1], 1.0);
}

// the configuration at one sample: Spline::impl_compute :353-372 + value :455-470 (the product accumulates the four
// terms in index order)
struct SplineQ {
  const double* base;
  const double* P;
  int nv;
  double b[4];
  __device__ __forceinline__ double operator()(int k) const {
    double v = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) v += P[(size_t)i * nv + k] * b[i];
    return base[k] + v;
  }
  __device__ __forceinline__ void freeflyer(int, double* out) const {  // (never reached: nq == nv is checked at launch)
#pragma unroll
    for (int k = 0; k < 7; ++k) out[k] = 0;
  }
};
__device__ __forceinline__ SplineQ splineQ(const double* base, const double* P, int nv, double s, double L) {
  SplineQ q;
  q.base = base; q.P = P; q.nv = nv;
  double u = 0;
  if (L != 0) {
    u = (s - 0.0) / L;
    if (u < 0.) u = 0.; else if (u > 1.) u = 1.;
  }
  splineBasisDerivative(0, u, q.b);
  return q;
}

// SolidSolidCollision::computeMaximalVelocity (src/continuous-validation/solid-solid-collision.cc:76-99) on the
// velocity bound of the spline over [ta, tb] (Spline::impl_velocityBound, src/path/spline.cc:423-434)
__device__ __forceinline__ double splineMaximalVelocity(const Tables& T, const DevBodyPair& B, const double* P, int nv, double L, double ta, double tb) {
  const double u0 = L == 0 ? 0 : (ta - 0.0) / L, u1 = L == 0 ? 0 : (tb - 0.0) / L;
  double ub[4];
  splineBasisBound(u0, u1, ub);
#pragma unroll
  for (int i = 0; i < 4; ++i) ub[i] /= L;
  auto sum = [&](const DevCoef* c, int b, int en) {
    double acc = 0;
    for (int x = b; x < en; ++x) {
      double s = 0;
      for (int k = 0; k < c[x].nv; ++k) {
        double vb = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) vb += fabs(P[(size_t)i * nv + c[x].idx_v + k]) * ub[i];
        s = fma(vb, vb, s);
      }
      acc += c[x].value * sqrt(s);
    }
    return acc;
  };
  const double ab = sum(T.coefs, B.coef_begin, B.coef_end);
  if (B.joint_b != 0) {
    const double ba = sum(T.rcoefs, B.rcoef_begin, B.rcoef_end);
    if (ba < ab) return ba;
  }
  return ab;
}

// BodyPairCollision::setupPath + collisionFreeInterval (src/continuous-validation/body-pair-collision.cc:101-173),
// refine_ = true: two rounds of "velocity bound on the interval just found -> new interval"
__device__ __forceinline__ double splineCollisionFreeInterval(const Tables& T, const DevBodyPair& B, const double* P, int nv, double L, double t,
                                                              double dlb, double& maxVelocity) {
  if (L == 0) {  // setupPath: maximalVelocity_ = inf, refine_ = false
    maxVelocity = CUDART_INF;
    return dlb / maxVelocity;
  }
  double Vm = splineMaximalVelocity(T, B, P, nv, L, 0.0, L);
  double Ti = dlb / Vm;
  maxVelocity = Vm;
  double tm = t - Ti, tM = t + Ti;
  const bool leftIsValid = tm < 0.0, rightIsValid = tM > L;
  if (leftIsValid && rightIsValid) return Ti;
  for (int i = 0; i < 4; ++i) {
    tm = t - (leftIsValid ? 0 : Ti);
    tM = t + (rightIsValid ? 0 : Ti);
    Vm = splineMaximalVelocity(T, B, P, nv, L, tm, tM);
    Ti = dlb / Vm;
  }
  maxVelocity = Vm;
  return Ti;
}

}  // namespace hppb200
