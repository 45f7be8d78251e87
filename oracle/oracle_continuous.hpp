// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_math.hpp header).
//
// CPU restatement of hpp-core's continuous path validation (Progressive / Dichotomy)
// for StraightPath edges.  Citations are relative to /root/reference.
// The Intervals restatement below IS pinned: tests/ check it against the reference's own
// src/continuous-validation/intervals.hh compiled into oracle/_ref (known answers of
// tests/test-intervals.cc + random sequences).  Everything numeric is PARITY UNPINNED.
#pragma once
#include <list>
#include <map>
#include <utility>

#include "oracle_validation.hpp"

namespace oracle {

typedef std::pair<double, double> interval_t;

// Sorted list of disjoint closed intervals; same observable behaviour as
// src/continuous-validation/intervals.hh:48-183 (touching intervals merge).
class Intervals {
 public:
  void clear() { v_.clear(); }
  void unionInterval(const interval_t& I) {
    // find first stored interval that ends at or after I.first (can merge on the left)
    size_t i = 0;
    while (i < v_.size() && v_[i].second < I.first) ++i;
    if (i == v_.size()) { v_.push_back(I); return; }
    if (I.second < v_[i].first) { v_.insert(v_.begin() + i, I); return; }
    double lo = I.first < v_[i].first ? I.first : v_[i].first;
    double hi = I.second;
    size_t j = i;
    while (j < v_.size() && v_[j].first <= I.second) {
      if (v_[j].second > hi) hi = v_[j].second;
      ++j;
    }
    v_[i] = interval_t(lo, hi);
    v_.erase(v_.begin() + i + 1, v_.begin() + j);
  }
  bool contains(const interval_t& I) const {
    for (const interval_t& s : v_)
      if (s.first <= I.first && I.second <= s.second) return true;
    return false;
  }
  const std::vector<interval_t>& list() const { return v_; }

 private:
  std::vector<interval_t> v_;
};

// src/continuous-validation/helper.hh:36-60
inline double first(const interval_t& I, bool reverse) { return reverse ? I.second : I.first; }
inline double second(const interval_t& I, bool reverse) { return first(I, !reverse); }

struct Coefficient {
  int joint;  // child joint whose velocity segment is used
  double value;
};

// src/continuous-validation/solid-solid-collision.cc + body-pair-collision.cc +
// include/hpp/core/continuous-validation/interval-validation.hh
struct SolidSolidCollision {
  const Device* dev = nullptr;
  int joint_a = 0, joint_b = 0;  // joint_b == 0: obstacles / universe body
  std::vector<CollisionPair> pairs;
  std::vector<int> pairIds;  // global geometry-pair index (for reports)
  std::vector<Coefficient> coefficients, coefficientsReverse;
  double tolerance = 0;
  // per-path state (IntervalValidation::path(), interval-validation.hh:84-90)
  bool valid_ = false;
  Intervals validInterval_;  // boost::icl::interval_set of closed intervals
  double maximalVelocity_ = 0;
  interval_t timeRange;
  const Path* path_ = nullptr;
  bool refine_ = false;

  // solid-solid-collision.cc:162-198 — climb from the larger index
  std::vector<int> computeSequenceOfJoints() const {
    std::vector<int> fromA, fromB;
    int ia = joint_a, ib = joint_b;
    while (ia != ib) {
      if (ia > ib) { fromA.push_back(ia); ia = dev->joints[ia].parent; }
      else { fromB.push_back(ib); ib = dev->joints[ib].parent; }
    }
    fromA.push_back(ia);
    std::vector<int> joints(fromA);
    joints.insert(joints.end(), fromB.rbegin(), fromB.rend());
    return joints;
  }
  // :200-232 — cumulativeLength starts at radius(body a) for BOTH directions (sic)
  std::vector<Coefficient> computeCoefficients(const std::vector<int>& joints, const std::vector<double>& radii) const {
    std::vector<Coefficient> coeff(joints.size() - 1);
    double cumulativeLength = radii[joint_a];
    for (size_t i = 0; i + 1 < joints.size(); ++i) {
      int child;
      if (dev->joints[joints[i]].parent == joints[i + 1]) child = joints[i];
      else if (dev->joints[joints[i + 1]].parent == joints[i]) child = joints[i + 1];
      else throw std::runtime_error("joint sequence is not a chain");
      double distance = dev->maximalDistanceToParent(child);
      coeff[i].joint = child;
      coeff[i].value = dev->upperBoundLinearVelocity(child) + cumulativeLength * dev->upperBoundAngularVelocity(child);
      cumulativeLength += distance;
    }
    return coeff;
  }
  // :234-244
  void setCoefficients(const std::vector<double>& radii) {
    std::vector<int> joints = computeSequenceOfJoints();
    coefficients = computeCoefficients(joints, radii);
    if (joint_b) {
      std::vector<int> rev(joints.rbegin(), joints.rend());
      coefficientsReverse = computeCoefficients(rev, radii);
    }
  }
  // :76-99
  double computeMaximalVelocity(const double* Vb) const {
    auto segNorm = [&](int j) {
      const Joint& J = dev->joints[j];
      double s = 0;
      for (int k = 0; k < J.nv; ++k) s = std::fma(Vb[J.idx_v + k], Vb[J.idx_v + k], s);
      return std::sqrt(s);
    };
    double ab = 0;
    for (const Coefficient& c : coefficients) ab += c.value * segNorm(c.joint);
    if (joint_b) {
      double ba = 0;
      for (const Coefficient& c : coefficientsReverse) ba += c.value * segNorm(c.joint);
      if (ba < ab) return ba;
    }
    return ab;
  }
  // body-pair-collision.cc:101-117 (StraightPath => refine_ = false)
  void setPath(const Path& path) {
    valid_ = false;
    validInterval_.clear();
    timeRange = interval_t(path.t0, path.t1);
    path_ = &path;
    refine_ = !path.straight();
    if (path.t1 - path.t0 == 0) {
      maximalVelocity_ = kInf;
      refine_ = false;
    } else {
      std::vector<double> Vb(dev->nv);
      path.velocityBound(Vb.data(), path.t0, path.t1);
      maximalVelocity_ = computeMaximalVelocity(Vb.data());
    }
  }
  // body-pair-collision.cc:119-173: two rounds of "velocity bound on the interval just found -> new interval"
  double collisionFreeInterval(double t, double dlb, double& maxVelocity) const {
    constexpr int Nrefine = 2;
    double T[1 + 2 * Nrefine], Vm[1 + 2 * Nrefine];
    Vm[0] = maxVelocity = maximalVelocity_;
    T[0] = dlb / maxVelocity;
    if (!refine_) return T[0];
    double tm = t - T[0], tM = t + T[0];
    const bool leftIsValid = tm < timeRange.first, rightIsValid = tM > timeRange.second;
    if (leftIsValid && rightIsValid) return T[0];
    std::vector<double> Vb(dev->nv);
    for (int i = 0; i < 2 * Nrefine; ++i) {
      tm = t - (leftIsValid ? 0 : T[i]);
      tM = t + (rightIsValid ? 0 : T[i]);
      path_->velocityBound(Vb.data(), tm, tM);
      Vm[i + 1] = computeMaximalVelocity(Vb.data());
      T[i + 1] = dlb / Vm[i + 1];
    }
    maxVelocity = Vm[2 * Nrefine];
    return T[2 * Nrefine];
  }
  // body-pair-collision.cc:175-198
  bool computeDistanceLowerBound(double& dlb, const DeviceData& data, int& hit, GjkStats* st) const {
    dlb = kInf;
    for (size_t i = 0; i < pairs.size(); ++i) {
      PairResult r = collidePair(*dev, data, pairs[i], st);
      if (r.collision) { hit = pairIds[i]; return false; }
      if (r.distanceLowerBound < dlb) dlb = r.distanceLowerBound;
    }
    return true;
  }
  // body-pair-collision.cc:45-99
  bool validateConfiguration(double t, interval_t& interval, const DeviceData& data, int& hit, GjkStats* st, uint64_t* evals) {
    if (valid_) { interval = timeRange; return true; }
    if (validInterval_.contains(interval)) return true;
    double dlb;
    if (evals) ++*evals;
    if (!computeDistanceLowerBound(dlb, data, hit, st)) return false;
    double halfLengthDist, halfLengthTol;
    if (dlb == kInf) { halfLengthDist = kInf; halfLengthTol = 0; }
    else {
      double Vm;
      halfLengthDist = collisionFreeInterval(t, dlb, Vm);  // (refine_ == false: dlb / maximalVelocity_, :119-128)
      halfLengthTol = (Vm != 0) ? 2 * tolerance / Vm : kInf;
    }
    interval.first = t - (halfLengthDist + halfLengthTol);
    interval.second = t + (halfLengthDist + halfLengthTol);
    validInterval_.unionInterval(interval);
    if (validInterval_.contains(timeRange)) valid_ = true;
    return true;
  }
};

enum OrderMode { ORDER_RESET_PER_EDGE = 0, ORDER_PERSISTENT = 1 };

// src/continuous-validation.cc
struct ContinuousValidation {
  const Device* dev;
  double tolerance;
  std::vector<SolidSolidCollision> intervalValidations_;  // construction order
  std::vector<SolidSolidCollision> pooled_;               // the pooled working copy
  std::vector<CollisionPair> allGeomPairs;                // flattened, for report indices
  int orderMode = ORDER_RESET_PER_EDGE;
  uint64_t maxIterations = 1u << 20;

  ContinuousValidation(const Device& d, double tol) : dev(&d), tolerance(tol) {
    std::vector<double> radii = d.bodyRadii();
    // Initialize::doExecute :63-98 — self pairs grouped per joint pair, first-appearance order
    std::map<std::pair<int, int>, size_t> bodyPairMap;
    for (auto& sp : d.selfPairs) {
      std::pair<int, int> jp(d.geoms[sp.first].joint, d.geoms[sp.second].joint);
      if (jp.first == jp.second) continue;
      auto it = bodyPairMap.find(jp);
      if (it == bodyPairMap.end()) {
        int joint1 = jp.first, joint2 = jp.second;
        if (joint2 == 0) std::swap(joint1, joint2);
        SolidSolidCollision ss;
        ss.dev = &d; ss.joint_a = joint2; ss.joint_b = joint1; ss.tolerance = tol;
        ss.setCoefficients(radii);
        intervalValidations_.push_back(ss);
        it = bodyPairMap.insert({jp, intervalValidations_.size() - 1}).first;
      }
      intervalValidations_[it->second].pairs.push_back({sp.first, sp.second, d.margin(jp.first, jp.second)});
    }
    // AddObstacle::doExecute :103-118 — one body pair per (joint with body, obstacle)
    for (size_t o = 0; o < d.obstacles.size(); ++o)
      for (int j = 1; j <= d.njoints(); ++j) {
        if (!d.isGlobal(o) || !d.jointHasBody(j)) continue;
        SolidSolidCollision ss;
        ss.dev = &d; ss.joint_a = j; ss.joint_b = 0; ss.tolerance = tol;
        for (size_t g = 0; g < d.geoms.size(); ++g)
          if (d.geoms[g].joint == j) ss.pairs.push_back({(int)g, -((int)o + 1), d.margin(j, 0)});
        ss.setCoefficients(radii);
        intervalValidations_.push_back(ss);
      }
    for (const PairOp& op : d.pairOps) apply(op);
    int id = 0;
    for (auto& ss : intervalValidations_)
      for (auto& p : ss.pairs) { ss.pairIds.push_back(id++); allGeomPairs.push_back(p); }
  }

  // ObstacleUserInterface calls on a ContinuousValidation (src/continuous-validation.cc:224-367)
  void apply(const PairOp& op) {
    const size_t n1 = dev->joints.size();
    switch (op.kind) {
      case HPPB200_OP_ADD_OBSTACLE_TO_JOINT: break;  // not part of the interface of a path validation
      case HPPB200_OP_ADD_OBSTACLE: {  // AddObstacle::doExecute :103-118, default request (margin 0)
        const std::vector<double> radii = dev->bodyRadii();
        for (int j = 1; j <= dev->njoints(); ++j) {
          if (!dev->jointHasBody(j)) continue;
          SolidSolidCollision ss;
          ss.dev = dev; ss.joint_a = j; ss.joint_b = 0; ss.tolerance = tolerance;
          for (size_t g = 0; g < dev->geoms.size(); ++g)
            if (dev->geoms[g].joint == j) ss.pairs.push_back({(int)g, -(op.a + 1), 0.0});
          ss.setCoefficients(radii);
          intervalValidations_.push_back(ss);
        }
        break;
      }
      case HPPB200_OP_REMOVE_OBSTACLE_FROM_JOINT: {  // :224-252 + SolidSolidCollision::removeObjectTo_b (:101-122)
        if (!dev->jointHasBody(op.b)) break;  // (the configuration validator only walks the inner objects of the body)
        bool removed = false;
        for (size_t i = 0; i < intervalValidations_.size();) {
          SolidSolidCollision& bpc = intervalValidations_[i];
          bool erased = false;
          if (bpc.joint_a == op.b && bpc.joint_b == 0) {
            std::vector<CollisionPair> kept;
            for (auto& p : bpc.pairs)
              if (p.second != -(op.a + 1)) kept.push_back(p);
            if (kept.size() != bpc.pairs.size()) {
              removed = true;
              bpc.pairs = kept;
              if (kept.empty()) { intervalValidations_.erase(intervalValidations_.begin() + i); erased = true; }
            }
          }
          if (!erased) ++i;
        }
        if (!removed) throw std::runtime_error("ContinuousValidation::removeObstacleFromJoint: obstacle is not registered as obstacle for joint");
        break;
      }
      case HPPB200_OP_FILTER_COLLISION_PAIRS:  // :266-300 (Parameterized treated as Constrained)
        for (size_t i = 0; i < intervalValidations_.size();) {
          const int m = op.relativeMotion[(size_t)intervalValidations_[i].joint_a * n1 + intervalValidations_[i].joint_b];
          if (m == 0 || m == 1) intervalValidations_.erase(intervalValidations_.begin() + i);
          else ++i;
        }
        break;
      case HPPB200_OP_SET_SECURITY_MARGINS:  // :302-330
        for (auto& bpc : intervalValidations_)
          for (auto& p : bpc.pairs) p.margin = op.matrix[(size_t)bpc.joint_a * n1 + bpc.joint_b];
        break;
      case HPPB200_OP_SET_MARGIN_BETWEEN_BODIES:  // :332-367 (every matching pair; `found` starts true: never throws)
        for (auto& bpc : intervalValidations_)
          for (auto& p : bpc.pairs)
            if ((p.first == op.a && p.second == op.b) || (p.first == op.b && p.second == op.a)) p.margin = op.value;
        break;
      default: throw std::invalid_argument("unknown pair op");
    }
  }

  // include/hpp/core/continuous-validation.hh:267-298 + src/continuous-validation.cc:128-148
  bool validateConfiguration(std::vector<SolidSolidCollision>& bpc, const double* q, double t, interval_t& interval,
                             int& hit, GjkStats* st, uint64_t* evals) const {
    interval.first = -kInf;
    interval.second = kInf;
    DeviceData data;
    dev->forwardKinematics(q, data);
    size_t smallest = 0;
    for (size_t i = 0; i < bpc.size(); ++i) {
      interval_t tmpInt = interval;
      if (!bpc[i].validateConfiguration(t, tmpInt, data, hit, st, evals)) return false;
      if (interval.second > tmpInt.second) smallest = i;
      interval.first = std::max(interval.first, tmpInt.first);
      interval.second = std::min(interval.second, tmpInt.second);
    }
    if (bpc.size() > 1 && smallest != 0) std::swap(bpc[0], bpc[smallest]);
    return true;
  }

  // src/continuous-validation.cc:193-205 (pool) + :216-222 (setPath)
  std::vector<SolidSolidCollision>& acquire(const Path& path) {
    if (orderMode == ORDER_RESET_PER_EDGE || pooled_.empty()) pooled_ = intervalValidations_;
    for (auto& b : pooled_) b.setPath(path);
    return pooled_;
  }
};

// src/continuous-validation/progressive.cc:78-129
struct Progressive : ContinuousValidation {
  Progressive(const Device& d, double tol) : ContinuousValidation(d, tol) {
    if (tol <= 0) throw std::runtime_error("tolerance should be positive for progressive continuous validation.");
  }
  PathReport validate(const Path& path, bool reverse, GjkStats* st = nullptr, uint64_t* evals = nullptr) {
    PathReport R;
    std::vector<SolidSolidCollision>& bpc = acquire(path);
    const interval_t tr(path.t0, path.t1);
    const double tmin = tr.first, tmax = tr.second;
    double lastValidTime = first(tr, reverse);
    double t = lastValidTime;
    unsigned finished = 0;
    bool valid = true;
    std::vector<double> q(dev->nq);
    interval_t interval;
    if (reverse ? t <= tmin : t >= tmax) finished++;
    while (finished < 2 && valid) {
      if (R.samples >= maxIterations) { R.valid = false; R.reason = HPPB200_REASON_ITER_CAP; R.parameter = t; R.lastValid = lastValidTime; return R; }
      path.eval(q.data(), t);
      double tprev = t;
      int hit = -1;
      R.samples++;
      if (!validateConfiguration(bpc, q.data(), t, interval, hit, st, evals)) {
        R.parameter = t; R.reason = HPPB200_REASON_COLLISION; R.index = hit;
        valid = false;
      } else {
        t = second(interval, reverse);
        lastValidTime = tprev;
      }
      if (reverse ? t <= tmin : t >= tmax) { t = reverse ? tmin : tmax; finished++; }
    }
    R.valid = valid;
    R.lastValid = valid ? (reverse ? tmin : tmax) : lastValidTime;
    return R;
  }
};

// src/continuous-validation/dichotomy.cc:68-132
struct Dichotomy : ContinuousValidation {
  Dichotomy(const Device& d, double tol) : ContinuousValidation(d, tol) {
    if (tol != 0) throw std::runtime_error("Dichotomy path validation method does not support penetration.");
  }
  PathReport validate(const Path& path, bool reverse, GjkStats* st = nullptr, uint64_t* evals = nullptr) {
    PathReport R;
    std::vector<SolidSolidCollision>& bpc = acquire(path);
    bool finished = false, valid = true;
    Intervals validSubset;
    const interval_t tr(path.t0, path.t1);
    double t = first(tr, reverse);
    validSubset.unionInterval(interval_t(t, t));
    std::vector<double> q(dev->nq);
    while (!finished) {
      if (R.samples >= maxIterations) { R.valid = false; R.reason = HPPB200_REASON_ITER_CAP; R.parameter = t;
        R.lastValid = reverse ? validSubset.list().back().first : validSubset.list().front().second; return R; }
      path.eval(q.data(), t);
      interval_t interval;
      int hit = -1;
      R.samples++;
      if (!validateConfiguration(bpc, q.data(), t, interval, hit, st, evals)) {
        R.parameter = t; R.reason = HPPB200_REASON_COLLISION; R.index = hit;
        valid = false;
      } else {
        validSubset.unionInterval(interval);
      }
      finished = (!valid || validSubset.contains(tr));
      const std::vector<interval_t>& L = validSubset.list();
      double t0 = second(reverse ? L.back() : L.front(), reverse);
      double t1 = second(tr, reverse);
      if (L.size() > 1) t1 = first(reverse ? L[L.size() - 2] : L[1], reverse);
      t = .5 * (t0 + t1);
    }
    R.valid = valid;
    if (!valid) R.lastValid = reverse ? validSubset.list().back().first : validSubset.list().front().second;
    else R.lastValid = reverse ? tr.first : tr.second;
    return R;
  }
};

}  // namespace oracle
