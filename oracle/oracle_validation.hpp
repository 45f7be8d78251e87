// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_math.hpp header).
//
// CPU restatement of hpp-core's configuration validation (one configuration at a
// time, first-hit semantics), StraightPath, and the discretized path validation.
// Each function cites the reference file:line it follows (paths relative to
// /root/reference).  PARITY UNPINNED for numeric outputs (reference tests assert only
// cross-method invariants); the invariants are re-asserted in tests/.
#pragma once
#include <algorithm>
#include <cstring>

#include "oracle_model.hpp"

namespace oracle {

// include/hpp/core/collision-pair.hh:45-81 — (first, second) objects + request margin.
// second < 0 encodes obstacle -(second+1).
struct CollisionPair {
  int first = 0, second = 0;
  double margin = 0;
};

struct PairResult {
  bool collision = false;
  double distanceLowerBound = kInf;
};

// CollisionPair::collide -> coal::ComputeCollision (Appendix A.4)
inline PairResult collidePair(const Device& dev, const DeviceData& data, const CollisionPair& p, GjkStats* st) {
  const Geometry& A = dev.geoms[p.first];
  const SE3& TA = data.oMg[p.first];
  const Geometry& B = p.second >= 0 ? dev.geoms[p.second] : dev.obstacles[-(p.second + 1)];
  const SE3& TB = p.second >= 0 ? data.oMg[p.second] : B.placement;
  double d = shapeDistance(A.shape, TA, B.shape, TB, st);
  PairResult r;
  r.distanceLowerBound = d - p.margin;
  r.collision = r.distanceLowerBound <= kCollisionThreshold;
  return r;
}

// src/obstacle-user.cc:53-76 (addObstacle: joint-major, inner objects of each body),
// :222-235 (addRobotCollisionPairs: active self pairs first), :166-196 (margins).
struct ObstacleUser {
  const Device* dev;
  std::vector<CollisionPair> cPairs;

  std::vector<CollisionPair> pPairs, dPairs;  // parameterized / disabled (filterCollisionPairs)

  explicit ObstacleUser(const Device& d) : dev(&d) {
    for (auto& sp : d.selfPairs) cPairs.push_back({sp.first, sp.second, 0.0});  // addRobotCollisionPairs
    for (size_t o = 0; o < d.obstacles.size(); ++o)
      if (d.isGlobal(o)) addObstacle((int)o);
    setSecurityMargins();
    for (const PairOp& op : d.pairOps) apply(op);
  }
  void addObstacle(int o) {
    // robot_->nbJoints()/jointAt(j): moving joints only, model joint j+1 (Appendix A.3)
    for (int j = 1; j <= dev->njoints(); ++j) addObstacleToJoint(o, j, false);
  }
  // src/obstacle-user.cc:60-76
  void addObstacleToJoint(int o, int joint, bool includeChildren) {
    for (size_t g = 0; g < dev->geoms.size(); ++g)
      if (dev->geoms[g].joint == joint) cPairs.push_back({(int)g, -(o + 1), 0.0});
    if (includeChildren)
      for (int c = 1; c <= dev->njoints(); ++c)
        if (dev->joints[c].parent == joint) addObstacleToJoint(o, c, includeChildren);
  }
  int jointOf(int ref) const { return ref >= 0 ? dev->geoms[ref].joint : 0; }
  void apply(const PairOp& op) {
    const size_t n1 = dev->joints.size();
    switch (op.kind) {
      case HPPB200_OP_ADD_OBSTACLE_TO_JOINT: addObstacleToJoint(op.a, op.b, op.flag != 0); break;
      case HPPB200_OP_ADD_OBSTACLE: addObstacle(op.a); break;  // cRequests_.push_back(defaultRequest_): margin 0
      case HPPB200_OP_REMOVE_OBSTACLE_FROM_JOINT:  // :87-118
        for (size_t g = 0; g < dev->geoms.size(); ++g) {
          if (dev->geoms[g].joint != op.b) continue;
          size_t nbDelPairs = 0;
          for (size_t i = 0; i < cPairs.size();) {
            if (cPairs[i].first == (int)g && cPairs[i].second == -(op.a + 1)) { cPairs.erase(cPairs.begin() + i); ++nbDelPairs; }
            else ++i;
          }
          if (nbDelPairs == 0) throw std::runtime_error("ObstacleUser::removeObstacleFromJoint: obstacle is not registered as obstacle for joint");
        }
        break;
      case HPPB200_OP_FILTER_COLLISION_PAIRS:  // :120-164
        for (size_t i = 0; i < cPairs.size();) {
          const int m = op.relativeMotion[(size_t)jointOf(cPairs[i].first) * n1 + jointOf(cPairs[i].second)];
          if (m == 1) { pPairs.push_back(cPairs[i]); cPairs.erase(cPairs.begin() + i); }        // Parameterized
          else if (m == 0) { dPairs.push_back(cPairs[i]); cPairs.erase(cPairs.begin() + i); }   // Constrained
          else ++i;
        }
        break;
      case HPPB200_OP_SET_SECURITY_MARGINS:  // :166-196
        for (auto* tab : {&cPairs, &pPairs, &dPairs})
          for (auto& p : *tab) p.margin = op.matrix[(size_t)jointOf(p.first) * n1 + jointOf(p.second)];
        break;
      case HPPB200_OP_SET_MARGIN_BETWEEN_BODIES: {  // :198-220
        auto setMargin = [&](std::vector<CollisionPair>& pairs) {
          for (auto& p : pairs)
            if ((p.first == op.a && p.second == op.b) || (p.first == op.b && p.second == op.a)) { p.margin = op.value; return true; }
          return false;
        };
        if (setMargin(cPairs) || setMargin(pPairs) || setMargin(dPairs)) break;
        throw std::invalid_argument("Could not find a collision pair between the two bodies");
      }
      default: throw std::invalid_argument("unknown pair op");
    }
  }
  void setSecurityMargins() {
    for (auto& p : cPairs) {
      int j1 = dev->geoms[p.first].joint;
      int j2 = p.second >= 0 ? dev->geoms[p.second].joint : 0;
      p.margin = dev->margin(j1, j2);
    }
  }
  // src/obstacle-user.cc:42-51 — sequential loop, stop at first hit.
  // the pairs CollisionValidation::validate walks, in order (what the engine's table holds)
  std::vector<CollisionPair> checkedPairs() const {
    std::vector<CollisionPair> v(cPairs);
    if (dev->checkParameterized) v.insert(v.end(), pPairs.begin(), pPairs.end());
    return v;
  }
  bool collide(const std::vector<CollisionPair>& pairs, const DeviceData& data, size_t& i, GjkStats* st) const {
    for (i = 0; i < pairs.size(); ++i)
      if (collidePair(*dev, data, pairs[i], st).collision) return true;
    return false;
  }
};

struct ConfigReport {
  int reason = HPPB200_REASON_VALID;
  int index = -1;  // first colliding pair / violated configuration rank
};

// src/collision-validation.cc:48-95
struct CollisionValidation : ObstacleUser {
  explicit CollisionValidation(const Device& d) : ObstacleUser(d) {}
  bool validate(const double* q, ConfigReport& rep, GjkStats* st = nullptr) const {
    DeviceData data;
    dev->forwardKinematics(q, data);
    size_t i = 0;
    bool hit = collide(cPairs, data, i, st);
    size_t base = 0;
    if (!hit && dev->checkParameterized) {  // :62-67
      hit = collide(pPairs, data, i, st);
      base = cPairs.size();  // reported as an index into cPairs ++ pPairs (the table the ABI exposes)
    }
    if (hit) {
      rep.reason = HPPB200_REASON_COLLISION;
      rep.index = (int)(base + i);
      return false;
    }
    rep = ConfigReport();
    return true;
  }
};

// src/joint-bound-validation.cc:45-80 — first violated rank wins.
struct JointBoundValidation {
  const Device* dev;
  explicit JointBoundValidation(const Device& d) : dev(&d) {}
  bool validate(const double* q, ConfigReport& rep) const {
    for (int i = 0; i < dev->nq; ++i) {
      // free-flyer quaternion ranks carry no bounds (lower > upper in pinocchio => skipped)
      if (dev->lower[i] > dev->upper[i]) continue;
      if (q[i] < dev->lower[i] || dev->upper[i] < q[i]) {
        rep.reason = HPPB200_REASON_JOINT_BOUND;
        rep.index = i;
        return false;
      }
    }
    rep = ConfigReport();
    return true;
  }
};

// src/config-validations.cc:40-49 — AND, first failure's report.
struct ConfigValidations {
  JointBoundValidation jb;
  CollisionValidation cv;
  bool checkBounds;
  ConfigValidations(const Device& d, bool bounds) : jb(d), cv(d), checkBounds(bounds) {}
  bool validate(const double* q, ConfigReport& rep, GjkStats* st = nullptr) const {
    if (checkBounds && !jb.validate(q, rep)) return false;
    return cv.validate(q, rep, st);
  }
};

// ---- Lie-group pieces (hpp-pinocchio LiegroupSpace, Appendix A.3) -------------
// q1 (-) q0, size nv.  R^n joints: plain difference; free-flyer: R^3 x SO(3) (oracle_math.hpp).
inline void lieDifference(const Device& dev, const double* q1, const double* q0, double* v) {
  for (int i = 1; i <= dev.njoints(); ++i) {
    const Joint& J = dev.joints[i];
    if (J.type == HPPB200_JOINT_FREEFLYER) ffDifference(q0 + J.idx_q, q1 + J.idx_q, v + J.idx_v);
    else v[J.idx_v] = q1[J.idx_q] - q0[J.idx_q];
  }
}
// interpolate(q0,q1,u) = q0 (+) u*(q1 (-) q0); R^n: fma(u, q1-q0, q0)
inline void lieInterpolate(const Device& dev, const double* q0, const double* q1, double u, double* q) {
  for (int i = 1; i <= dev.njoints(); ++i) {
    const Joint& J = dev.joints[i];
    if (J.type == HPPB200_JOINT_FREEFLYER) ffInterpolate(q0 + J.idx_q, q1 + J.idx_q, u, q + J.idx_q);
    else q[J.idx_q] = std::fma(u, q1[J.idx_q] - q0[J.idx_q], q0[J.idx_q]);
  }
}

// src/weighed-distance.cc:272-289 — sqrt(sum_j squaredDistance_j(q1,q2,w_j^2)).
// Vector-space rotation joints: w^2 * dq^2; translations unweighted (Appendix A.3).
inline double weighedDistance(const Device& dev, const double* q1, const double* q2) {
  double res = 0;
  for (int i = 1; i <= dev.njoints(); ++i) {
    const Joint& J = dev.joints[i];
    double w2 = dev.weights[i - 1] * dev.weights[i - 1];
    double d;
    if (J.type == HPPB200_JOINT_FREEFLYER) d = ffSquaredDistance(q1 + J.idx_q, q2 + J.idx_q, w2);
    else {
      double dq = q2[J.idx_q] - q1[J.idx_q];
      d = (J.type <= HPPB200_JOINT_REVOLUTE_Z) ? w2 * (dq * dq) : dq * dq;
    }
    res += d;
  }
  return std::sqrt(res);
}

// include/hpp/core/path.hh: what the validation algorithms need of a path
struct Path {
  const Device* dev = nullptr;
  double t0 = 0, t1 = 0;  // paramRange == timeRange (no time parameterization)
  virtual ~Path() {}
  virtual void eval(double* q, double param) const = 0;
  // Path::velocityBound(result, t0, t1) on a sub-interval (impl_velocityBound)
  virtual void velocityBound(double* vb, double ta, double tb) const = 0;
  // BodyPairCollision::setupPath (body-pair-collision.cc:101-105): refine_ = path is not a StraightPath
  virtual bool straight() const = 0;
  double length() const { return t1 - t0; }
};

// src/straight-path.cc:91-158, src/steering-method/straight.cc:38-51
struct StraightPath : Path {
  std::vector<double> initial, end;
  StraightPath(const Device& d, const double* q1, const double* q2, double L) : initial(q1, q1 + d.nq), end(q2, q2 + d.nq) {
    dev = &d; t0 = 0; t1 = L;
  }
  bool straight() const override { return true; }
  void velocityBound(double* vb, double, double) const override { velocityBound(vb); }
  // impl_compute :91-110 — endpoints returned bit-exactly
  void eval(double* q, double param) const override {
    const double L = t1 - t0;
    if (param == t0 || L == 0) { std::memcpy(q, initial.data(), sizeof(double) * dev->nq); return; }
    if (param == t1) { std::memcpy(q, end.data(), sizeof(double) * dev->nq); return; }
    double u = (param - t0) / L;
    lieInterpolate(*dev, initial.data(), end.data(), u, q);
  }
  // impl_velocityBound :132-140 — |end (-) initial| / L, zero for zero-length paths
  void velocityBound(double* vb) const {
    if (t0 == t1) { for (int i = 0; i < dev->nv; ++i) vb[i] = 0; return; }
    lieDifference(*dev, end.data(), initial.data(), vb);
    const double L = t1 - t0;
    for (int i = 0; i < dev->nv; ++i) vb[i] = std::fabs(vb[i]) / L;
  }
};

// src/path/spline.cc, include/hpp/core/path/spline.hh: path::Spline<BernsteinBasis, 3>, the path type of the reference's
// non-straight continuous-validation test (tests/test-continuous-validation.cc:288-427).  q(t) = base (+) P^T b(u),
// u = (t - t0) / (t1 - t0) clamped to [0, 1], b the four cubic Bernstein polynomials, P [4][nv] the parameters
// (tangent vectors at `base`).  Vector-space configuration spaces only (nq == nv: the (+) is a sum).
struct SplinePath : Path {
  std::vector<double> base, P;  // P row-major [4][nv]
  SplinePath(const Device& d, const double* base_, const double* params, double L) : base(base_, base_ + d.nq), P(params, params + 4 * d.nv) {
    if (d.nq != d.nv) throw std::runtime_error("spline paths: vector-space configuration spaces only");
    dev = &d; t0 = 0; t1 = L;
  }
  bool straight() const override { return false; }
  // spline_basis_function<BernsteinBasis, 3>::derivative (src/path/spline.cc:131-159), order k of the basis at t
  static void basisDerivative(int k, double t, double res[4]) {
    const int Degree = 3;
    static const double factors[4] = {1, 1, 2, 6};
    static const int binom[4][4] = {{1, 0, 0, 0}, {1, 1, 0, 0}, {1, 2, 1, 0}, {1, 3, 3, 1}};
    for (int i = 0; i < 4; ++i) res[i] = 0;
    if (k > Degree) return;
    double pT[4], pO[4];
    pT[0] = 1; pO[0] = 1;
    const double oneMinusT = 1 - t;
    for (int i = 1; i < 4; ++i) { pT[i] = pT[i - 1] * t; pO[i] = pO[i - 1] * oneMinusT; }
    for (int i = 0; i < 4; ++i)
      for (int p = std::max(0, k + i - Degree); p <= std::min(i, k); ++p) {
        const int ip = i - p;
        res[i] += double(((k - p) % 2 == 0 ? 1 : -1) * binom[k][p]) * pT[ip] * pO[Degree - k - ip] / double(factors[ip] * factors[Degree - k - ip]);
      }
    for (int i = 0; i < 4; ++i) res[i] *= factors[Degree];
  }
  // impl_compute :353-372 and value :455-470 (the matrix product accumulates the four terms in index order)
  void eval(double* q, double s) const override {
    double u = 0;
    if (t1 - t0 != 0) {
      u = (s - t0) / (t1 - t0);
      if (u < 0.) u = 0.; else if (u > 1.) u = 1.;
    }
    double b[4];
    basisDerivative(0, u, b);
    const int nv = dev->nv;
    for (int k = 0; k < nv; ++k) {
      double v = 0;
      for (int i = 0; i < 4; ++i) v += P[(size_t)i * nv + k] * b[i];
      q[k] = base[k] + v;
    }
  }
  // spline_basis_function<BernsteinBasis, 3>::bound (src/path/spline.cc:204-247), Degree = 3: b_up(1) = b_up(2) = 1
  static void basisBound(double u0, double u1, double ub[4]) {
    const int Degree = 3;
    double b0[4], b1[4];
    basisDerivative(1, u0, b0);
    basisDerivative(1, u1, b1);
    for (int i = 0; i < 4; ++i) ub[i] = std::max(std::fabs(b0[i]), std::fabs(b1[i]));
    if (u0 * Degree < 2 && 2 < u1 * Degree) ub[1] = std::max(ub[1], 1.0);
    if (u0 * Degree < Degree - 2 && Degree - 2 < u1 * Degree) ub[Degree - 1] = std::max(ub[Degree - 1], 1.0);
  }
  // impl_velocityBound :423-434
  void velocityBound(double* vb, double ta, double tb) const override {
    const double len = t1 - t0;
    const double u0 = len == 0 ? 0 : (ta - t0) / len, u1 = len == 0 ? 0 : (tb - t0) / len;
    double ub[4];
    basisBound(u0, u1, ub);
    for (int i = 0; i < 4; ++i) ub[i] /= len;
    const int nv = dev->nv;
    for (int k = 0; k < nv; ++k) {
      double v = 0;
      for (int i = 0; i < 4; ++i) v += std::fabs(P[(size_t)i * nv + k]) * ub[i];
      vb[k] = v;
    }
  }
};

struct PathReport {
  bool valid = true;
  double parameter = std::numeric_limits<double>::quiet_NaN();  // report->parameter
  double lastValid = 0;  // forward: validPart = extract(tmin,lastValid); reverse: extract(lastValid,tmax)
  int reason = HPPB200_REASON_VALID;
  int index = -1;
  uint64_t samples = 0;
};

// src/path-validation/discretized.cc:55-107
struct Discretized {
  ConfigValidations cfg;
  double stepSize;
  Discretized(const Device& d, double step, bool bounds) : cfg(d, bounds), stepSize(step) {}
  PathReport validate(const Path& path, bool reverse, GjkStats* st = nullptr) const {
    PathReport R;
    const double tmin = path.t0, tmax = path.t1;
    unsigned finished = 0;
    std::vector<double> q(path.dev->nq);
    double T1, step, lastValidTime;
    if (reverse) { lastValidTime = tmax; T1 = tmin; step = -stepSize; }
    else { lastValidTime = tmin; T1 = tmax; step = stepSize; }
    double t = lastValidTime;
    bool valid = true;
    while (finished < 2 && valid) {
      path.eval(q.data(), t);
      ConfigReport cr;
      R.samples++;
      if (!cfg.validate(q.data(), cr, st)) {
        R.parameter = t; R.reason = cr.reason; R.index = cr.index;
        valid = false;
      } else {
        lastValidTime = t;
        t += step;
      }
      if ((reverse && t < T1) || (!reverse && t > T1)) { t = T1; finished++; }
    }
    R.valid = valid;
    R.lastValid = valid ? (reverse ? tmin : tmax) : lastValidTime;
    return R;
  }
};

}  // namespace oracle
