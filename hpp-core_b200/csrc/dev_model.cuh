// Device view of the compiled model blob + TMA staging into shared memory.
// Table layouts are defined in engine.hpp (host) and mirrored here.
#pragma once
#include "dev_narrow.cuh"
#include "engine.hpp"

namespace hppb200 {

static_assert(sizeof(DevShape) == sizeof(DevShapeH), "shape record mismatch");

// Typed pointers into a staged (shared-memory) or global copy of the blob.
struct Tables {
  const DevJoint* joints;
  const DevGeom* geoms;
  const DevObstacle* obstacles;
  const DevPair* cpairs;
  const DevBodyPair* bodypairs;
  const DevPair* bgpairs;
  const DevCoef* coefs;
  const DevCoef* rcoefs;
  const double* lower;
  const double* upper;
  const double* weights;
};

__device__ __forceinline__ Tables makeTables(const unsigned char* base, const ModelView& mv) {
  Tables t;
  t.joints = (const DevJoint*)(base + mv.off_joints);
  t.geoms = (const DevGeom*)(base + mv.off_geoms);
  t.obstacles = (const DevObstacle*)(base + mv.off_obstacles);
  t.cpairs = (const DevPair*)(base + mv.off_cpairs);
  t.bodypairs = (const DevBodyPair*)(base + mv.off_bodypairs);
  t.bgpairs = (const DevPair*)(base + mv.off_bgpairs);
  t.coefs = (const DevCoef*)(base + mv.off_coefs);
  t.rcoefs = (const DevCoef*)(base + mv.off_rcoefs);
  t.lower = (const double*)(base + mv.off_lower);
  t.upper = (const double*)(base + mv.off_upper);
  t.weights = (const double*)(base + mv.off_weights);
  return t;
}

// Stage `bytes` (multiple of 16, both pointers 16-byte aligned) from global to shared memory
// with one TMA bulk copy (cp.async.bulk, SASS UBLKCP) completed through an mbarrier.
// Must be called by every thread of the CTA; returns when the data is visible to all.
__device__ __forceinline__ void stageBlobTMA(unsigned char* sdst, const unsigned char* gsrc, uint32_t bytes,
                                             uint64_t* bar) {
  const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(bar);
  const uint32_t dst_s = (uint32_t)__cvta_generic_to_shared(sdst);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(bytes) : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s),
        "l"(gsrc), "r"(bytes), "r"(bar_s)
        : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar_s), "r"(0u)
        : "memory");
  }
}

// Configuration accessors handed to the FK: entry k of a vector-space joint, or the 7 entries of a free-flyer.
struct ConfigQ {  // a stored configuration
  const double* q;
  __device__ __forceinline__ double operator()(int k) const { return q[k]; }
  __device__ __forceinline__ void freeflyer(int idx_q, double* out) const {
#pragma unroll
    for (int k = 0; k < 7; ++k) out[k] = q[idx_q + k];
  }
};
// Path::eval -> StraightPath::impl_compute (src/straight-path.cc:91-110): qmode 0 / 1 return the stored end
// points bit-exactly, 2 is LiegroupSpace::interpolate(q1, q2, u)
struct EdgeQ {
  const double* q1;
  const double* q2;
  int qmode;
  double u;
  __device__ __forceinline__ double operator()(int k) const {
    const double a = q1[k];
    if (qmode == 0) return a;
    const double b = q2[k];
    if (qmode == 1) return b;
    return fma(u, b - a, a);  // R^n
  }
  __device__ __forceinline__ void freeflyer(int idx_q, double* out) const {
    if (qmode == 2) { ffInterpolate(q1 + idx_q, q2 + idx_q, u, out); return; }
    const double* s = (qmode == 0 ? q1 : q2) + idx_q;
#pragma unroll
    for (int k = 0; k < 7; ++k) out[k] = s[k];
  }
};
__device__ __forceinline__ EdgeQ edgeQ(const double* q1, const double* q2, double t, double L) {
  EdgeQ e;
  e.q1 = q1; e.q2 = q2; e.u = 0;
  if (t == 0.0 || L == 0.0) e.qmode = 0;
  else if (t == L) e.qmode = 1;
  else { e.qmode = 2; e.u = (t - 0.0) / L; }
  return e;
}

// |Vb| over the velocity segment of one joint, Vb = |q2 (-) q1| / L (StraightPath::impl_velocityBound,
// src/straight-path.cc:132-140; the segment norm of SolidSolidCollision::computeMaximalVelocity :76-99)
__device__ __forceinline__ double jointVelocityNorm(const double* q1, const double* q2, int idx_q, int nvj, double L) {
  if (nvj == 1) {
    const double vb = fabs(q2[idx_q] - q1[idx_q]) / L;
    return sqrt(fma(vb, vb, 0.0));
  }
  double v[6];
  ffDifference(q1 + idx_q, q2 + idx_q, v);
  double s = 0;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const double vb = fabs(v[k]) / L;
    s = fma(vb, vb, s);
  }
  return sqrt(s);
}

// pinocchio joint transform composed with the joint placement: liMi = P * M_joint(q)
template <class Q>
__device__ __forceinline__ T3 jointLocal(const DevJoint& J, const Q& q) {
  T3 P = load12(J.P);
  T3 M = identityT();
  if (J.type <= HPPB200_JOINT_REVOLUTE_Z) {
    double s, c;
    sincos_portable(q(J.idx_q), &s, &c);
    // static indices only (keeps M in registers)
    if (J.type == HPPB200_JOINT_REVOLUTE_X) { M.r[4] = c; M.r[5] = -s; M.r[7] = s; M.r[8] = c; }
    else if (J.type == HPPB200_JOINT_REVOLUTE_Y) { M.r[8] = c; M.r[6] = -s; M.r[2] = s; M.r[0] = c; }
    else { M.r[0] = c; M.r[1] = -s; M.r[3] = s; M.r[4] = c; }
  } else if (J.type <= HPPB200_JOINT_PRISMATIC_Z) {
    const int a = J.type - HPPB200_JOINT_PRISMATIC_X;
    double v = q(J.idx_q);
    M.t = mk(a == 0 ? v : 0.0, a == 1 ? v : 0.0, a == 2 ? v : 0.0);
  } else {  // free-flyer: translation + unit quaternion (x y z w), Eigen toRotationMatrix operation order
    double f[7];
    q.freeflyer(J.idx_q, f);
    M.t = mk(f[0], f[1], f[2]);
    const double x = f[3], y = f[4], z = f[5], w = f[6];
    const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w;
    const double txx = tx * x, txy = ty * x, txz = tz * x;
    const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
    M.r[0] = 1 - (tyy + tzz); M.r[1] = txy - twz; M.r[2] = txz + twy;
    M.r[3] = txy + twz; M.r[4] = 1 - (txx + tzz); M.r[5] = tyz - twx;
    M.r[6] = txz - twy; M.r[7] = tyz + twx; M.r[8] = 1 - (txx + tyy);
  }
  return compose(P, M);
}


// FK for one configuration (pinocchio::forwardKinematics + updateGeometryPlacements):
//   liMi = placement * M_joint(q); oMi = oMi[parent] * liMi; oMg = oMi[joint] * placement_g.
// `store` is SoA: element k of slot s for this thread at store[(s*12+k)*stride].
// Geometry g uses slot g; for tree-shaped robots joint j (1-based) uses slot ngeoms + j - 1.
// `q(k)` returns configuration entry k.
template <bool TREE, class Q>
__device__ __forceinline__ void forwardKinematics(const Tables& T, const ModelView& mv, const Q& q, double* store,
                                                  size_t stride) {
  for (int g = 0; g < mv.ngeoms; ++g)
    if (T.geoms[g].joint == 0) store12s(store + (size_t)g * 12 * stride, stride, load12(T.geoms[g].P));
  T3 cur = identityT();
  for (int j = 1; j <= mv.njoints; ++j) {
    const DevJoint& J = T.joints[j - 1];
    T3 li = jointLocal(J, q);
    if (J.parent == 0) cur = li;
    else if (!TREE || J.parent == j - 1) cur = compose(cur, li);
    else cur = compose(load12s(store + (size_t)(mv.ngeoms + J.parent - 1) * 12 * stride, stride), li);
    if (TREE) store12s(store + (size_t)(mv.ngeoms + j - 1) * 12 * stride, stride, cur);
    for (int g = 0; g < mv.ngeoms; ++g)
      if (T.geoms[g].joint == j) store12s(store + (size_t)g * 12 * stride, stride, compose(cur, load12(T.geoms[g].P)));
  }
}

// distance_lower_bound (distance - security margin) of one canonically ordered pair
__device__ __forceinline__ double pairLowerBound(const Tables& T, const DevPair& pr, const double* store, size_t stride,
                                                 const double* __restrict__ verts, GjkCounters& cnt) {
  const DevShape* sa;
  const DevShape* sb;
  T3 TA, TB;
  if (pr.a >= 0) { sa = (const DevShape*)&T.geoms[pr.a].shape; TA = load12s(store + (size_t)pr.a * 12 * stride, stride); }
  else { const DevObstacle& O = T.obstacles[-(pr.a + 1)]; sa = (const DevShape*)&O.shape; TA = load12(O.P); }
  if (pr.b >= 0) { sb = (const DevShape*)&T.geoms[pr.b].shape; TB = load12s(store + (size_t)pr.b * 12 * stride, stride); }
  else { const DevObstacle& O = T.obstacles[-(pr.b + 1)]; sb = (const DevShape*)&O.shape; TB = load12(O.P); }
  return shapeDistance(pr.kind, *sa, TA, *sb, TB, verts, cnt) - pr.margin;
}

static inline bool modelIsTree(const Model* m) {
  for (int k = 0; k < m->njoints; ++k)
    if (m->joints[k].parent != k) return true;  // joint k+1 whose parent is not joint k
  return false;
}

}  // namespace hppb200
