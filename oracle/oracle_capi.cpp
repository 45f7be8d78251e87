// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_math.hpp header).
// extern "C" surface of the oracle for ctypes (tests/, smoke(), bench.py cpu_baseline).
// Batch drivers use OpenMP over independent configurations / edges, one working copy of
// the body-pair state per thread, as the reference does with its pools
// (src/continuous-validation.cc:193-205, tests/test-continuous-validation.cc:324).
#include <omp.h>

#include <memory>

#include "oracle_continuous.hpp"
#include "oracle_nn.hpp"

using namespace oracle;

struct OracleHandle {
  std::unique_ptr<Device> dev;
  std::unique_ptr<Roadmap> roadmap;
  GjkStats stats;
  uint64_t pairEvals = 0, samples = 0;
};

static thread_local std::string g_err;

extern "C" {

const char* oracle_last_error() { return g_err.c_str(); }

int oracle_create(const hppb200_model_desc* desc, OracleHandle** out) {
  try {
    auto h = new OracleHandle;
    h->dev.reset(new Device(*desc));
    h->roadmap.reset(new Roadmap(*h->dev));
    *out = h;
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}
void oracle_destroy(OracleHandle* h) { delete h; }
int oracle_nq(OracleHandle* h) { return h->dev->nq; }
int oracle_nv(OracleHandle* h) { return h->dev->nv; }

void oracle_sincos(const double* x, size_t n, double* s, double* c) {
  for (size_t i = 0; i < n; ++i) sincos_portable(x[i], s + i, c + i);
}

void oracle_atan(const double* x, size_t n, double* y) {
  for (size_t i = 0; i < n; ++i) y[i] = atan_portable(x[i]);
}
// StraightPath::eval at parameter t of the edge (q1, q2, L), and q2 (-) q1
void oracle_interpolate(OracleHandle* h, const double* q1, const double* q2, const double* L, const double* t, size_t n, double* q) {
  const Device& d = *h->dev;
  for (size_t i = 0; i < n; ++i) StraightPath(d, q1 + i * d.nq, q2 + i * d.nq, L[i]).eval(q + i * d.nq, t[i]);
}
void oracle_difference(OracleHandle* h, const double* q1, const double* q2, size_t n, double* v) {
  const Device& d = *h->dev;
  for (size_t i = 0; i < n; ++i) lieDifference(d, q2 + i * d.nq, q1 + i * d.nq, v + i * d.nv);
}

// ---- tables -----------------------------------------------------------------
int oracle_num_config_pairs(OracleHandle* h) { return (int)ObstacleUser(*h->dev).checkedPairs().size(); }
void oracle_get_config_pairs(OracleHandle* h, int32_t* a, int32_t* b, double* margin) {
  const std::vector<CollisionPair> cp = ObstacleUser(*h->dev).checkedPairs();
  for (size_t i = 0; i < cp.size(); ++i) { a[i] = cp[i].first; b[i] = cp[i].second; margin[i] = cp[i].margin; }
}
int oracle_num_body_pairs(OracleHandle* h) { return (int)ContinuousValidation(*h->dev, 0).intervalValidations_.size(); }
int oracle_num_body_geom_pairs(OracleHandle* h) { return (int)ContinuousValidation(*h->dev, 0).allGeomPairs.size(); }
int oracle_num_coefficients(OracleHandle* h, int reverse) {
  ContinuousValidation cv(*h->dev, 0);
  size_t n = 0;
  for (auto& s : cv.intervalValidations_) n += reverse ? s.coefficientsReverse.size() : s.coefficients.size();
  return (int)n;
}
void oracle_get_body_pairs(OracleHandle* h, int32_t* ja, int32_t* jb, int32_t* pair_begin, int32_t* coef_begin, int32_t* rcoef_begin) {
  ContinuousValidation cv(*h->dev, 0);
  int pb = 0, cb = 0, rb = 0;
  size_t i = 0;
  for (auto& s : cv.intervalValidations_) {
    ja[i] = s.joint_a; jb[i] = s.joint_b; pair_begin[i] = pb; coef_begin[i] = cb; rcoef_begin[i] = rb;
    pb += (int)s.pairs.size(); cb += (int)s.coefficients.size(); rb += (int)s.coefficientsReverse.size();
    ++i;
  }
  pair_begin[i] = pb; coef_begin[i] = cb; rcoef_begin[i] = rb;
}
void oracle_get_body_geom_pairs(OracleHandle* h, int32_t* a, int32_t* b, double* margin) {
  ContinuousValidation cv(*h->dev, 0);
  for (size_t i = 0; i < cv.allGeomPairs.size(); ++i) { a[i] = cv.allGeomPairs[i].first; b[i] = cv.allGeomPairs[i].second; margin[i] = cv.allGeomPairs[i].margin; }
}
void oracle_get_coefficients(OracleHandle* h, int reverse, int32_t* joint, double* value) {
  ContinuousValidation cv(*h->dev, 0);
  size_t k = 0;
  for (auto& s : cv.intervalValidations_)
    for (auto& c : (reverse ? s.coefficientsReverse : s.coefficients)) { joint[k] = c.joint; value[k] = c.value; ++k; }
}
void oracle_get_body_radii(OracleHandle* h, double* r) {
  std::vector<double> v = h->dev->bodyRadii();
  for (size_t i = 0; i < v.size(); ++i) r[i] = v[i];
}

// ---- FK / distances ------------------------------------------------------------
void oracle_geometry_placements(OracleHandle* h, const double* q, size_t n, double* oMg) {
  const Device& d = *h->dev;
  DeviceData data;
  for (size_t i = 0; i < n; ++i) {
    d.forwardKinematics(q + i * d.nq, data);
    for (size_t g = 0; g < d.geoms.size(); ++g) data.oMg[g].to12(oMg + (i * d.geoms.size() + g) * 12);
  }
}
void oracle_pair_distances(OracleHandle* h, const double* q, size_t n, double* dist) {
  const Device& d = *h->dev;
  const std::vector<CollisionPair> cp = ObstacleUser(d).checkedPairs();
  DeviceData data;
  for (size_t i = 0; i < n; ++i) {
    d.forwardKinematics(q + i * d.nq, data);
    for (size_t p = 0; p < cp.size(); ++p)
      dist[i * cp.size() + p] = collidePair(d, data, cp[p], nullptr).distanceLowerBound;
  }
}
// CollisionValidation::validate with computeAllContacts_ (src/collision-validation.cc:72-89): the first colliding pair,
// then every remaining pair of the same table that collides.  mask over the table cPairs ++ pPairs.
void oracle_all_collisions(OracleHandle* h, const double* q, size_t n, uint8_t* valid, int32_t* first, uint32_t* mask, int32_t* count) {
  const Device& d = *h->dev;
  ObstacleUser ou(d);
  const size_t np = ou.checkedPairs().size(), words = (np + 31) / 32;
  DeviceData data;
  for (size_t i = 0; i < n; ++i) {
    d.forwardKinematics(q + i * d.nq, data);
    for (size_t w = 0; w < words; ++w) mask[i * words + w] = 0u;
    const std::vector<CollisionPair>* pairs = &ou.cPairs;
    size_t base = 0, iPair = 0;
    bool collide = ou.collide(ou.cPairs, data, iPair, nullptr);
    if (!collide && d.checkParameterized) {
      collide = ou.collide(ou.pPairs, data, iPair, nullptr);
      pairs = &ou.pPairs;
      base = ou.cPairs.size();
    }
    int c = 0;
    if (collide) {
      first[i] = (int32_t)(base + iPair);
      mask[i * words + ((base + iPair) >> 5)] |= 1u << ((base + iPair) & 31);
      ++c;
      for (++iPair; iPair < pairs->size(); ++iPair)
        if (collidePair(d, data, (*pairs)[iPair], nullptr).collision) { mask[i * words + ((base + iPair) >> 5)] |= 1u << ((base + iPair) & 31); ++c; }
    } else first[i] = -1;
    valid[i] = collide ? 0 : 1;
    count[i] = c;
  }
}
// DistanceBetweenObjects (src/distance-between-objects.cc:41-77): obstacle-major (inner object, obstacle) pairs, plain distance
int oracle_num_distance_pairs(OracleHandle* h) {
  const Device& d = *h->dev;
  int c = 0;
  for (size_t o = 0; o < d.obstacles.size(); ++o)
    for (int j = 1; j <= d.njoints(); ++j)
      for (size_t g = 0; g < d.geoms.size(); ++g) c += d.geoms[g].joint == j;
  return c;
}
void oracle_distance_between_objects(OracleHandle* h, const double* q, size_t n, double* dist) {
  const Device& d = *h->dev;
  const size_t per = (size_t)oracle_num_distance_pairs(h);
  DeviceData data;
  for (size_t i = 0; i < n; ++i) {
    d.forwardKinematics(q + i * d.nq, data);
    size_t k = 0;
    for (size_t o = 0; o < d.obstacles.size(); ++o)
      for (int j = 1; j <= d.njoints(); ++j)
        for (size_t g = 0; g < d.geoms.size(); ++g)
          if (d.geoms[g].joint == j) dist[i * per + k++] = collidePair(d, data, CollisionPair{(int)g, -((int)o + 1), 0.0}, nullptr).distanceLowerBound;
  }
}
// distance between two shapes given as (type, params[4], placement[12]); convex verts optional
double oracle_shape_distance(int ta, const double* pa, const double* Ta, const double* va, int na, int tb,
                             const double* pb, const double* Tb, const double* vb, int nb) {
  Shape A, B;
  A.type = ta; B.type = tb;
  for (int k = 0; k < 4; ++k) { A.p[k] = pa[k]; B.p[k] = pb[k]; }
  A.verts = va; A.nverts = na; B.verts = vb; B.nverts = nb;
  return shapeDistance(A, SE3::from12(Ta), B, SE3::from12(Tb), nullptr);
}

// ---- configuration validation ----------------------------------------------------
int oracle_validate_configs(OracleHandle* h, const double* q, size_t n, uint32_t flags, uint8_t* valid,
                            int32_t* first_pair, uint8_t* reason, int nthreads) {
  const Device& d = *h->dev;
  ConfigValidations cv(d, flags & HPPB200_CHECK_JOINT_BOUNDS);
  uint64_t calls = 0, iters = 0, closed = 0;
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads) reduction(+ : calls, iters, closed)
  for (long i = 0; i < (long)n; ++i) {
    ConfigReport rep;
    GjkStats st;
    bool ok = cv.validate(q + (size_t)i * d.nq, rep, &st);
    valid[i] = ok ? 1 : 0;
    if (first_pair) first_pair[i] = rep.index;
    if (reason) reason[i] = (uint8_t)rep.reason;
    calls += st.calls; iters += st.iterations; closed += st.closed;
  }
  h->stats.calls += calls; h->stats.iterations += iters; h->stats.closed += closed;
  return 0;
}

// ---- edges -----------------------------------------------------------------------
void oracle_straight_lengths(OracleHandle* h, const double* q1, const double* q2, size_t n, double* L) {
  const Device& d = *h->dev;
  for (size_t i = 0; i < n; ++i) L[i] = weighedDistance(d, q1 + i * d.nq, q2 + i * d.nq);
}

}  // extern "C"

// the edge loop of oracle_validate_edges, for any path type: makePath(i) builds edge i's path
template <typename MakePath>
static int validatePathsImpl(OracleHandle* h, int method, double param, int reverse, size_t n, uint32_t flags, uint8_t* valid, double* t_fail,
                             double* t_last_valid, int32_t* pair, uint8_t* reason, uint64_t* samples, int order_mode, uint64_t max_iter,
                             int nthreads, MakePath makePath) {
  const Device& d = *h->dev;
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  if (order_mode == ORDER_PERSISTENT) nthreads = 1;  // history-dependent: sequential like one reference thread
  int err = 0;
  uint64_t evals = 0, nsamples = 0, calls = 0, iters = 0, closed = 0;
#pragma omp parallel num_threads(nthreads) reduction(+ : evals, nsamples, calls, iters, closed)
  {
    GjkStats st;  // per thread: GJK calls / iterations / closed-form calls of the edges it validates
    try {
      std::unique_ptr<Discretized> disc;
      std::unique_ptr<Progressive> prog;
      std::unique_ptr<Dichotomy> dich;
      if (method == HPPB200_METHOD_DISCRETIZED) disc.reset(new Discretized(d, param, flags & HPPB200_CHECK_JOINT_BOUNDS));
      else if (method == HPPB200_METHOD_PROGRESSIVE) { prog.reset(new Progressive(d, param)); prog->orderMode = order_mode; if (max_iter) prog->maxIterations = max_iter; }
      else { dich.reset(new Dichotomy(d, param)); dich->orderMode = order_mode; if (max_iter) dich->maxIterations = max_iter; }
#pragma omp for schedule(dynamic, 16)
      for (long i = 0; i < (long)n; ++i) {
        const auto path = makePath((size_t)i);
        PathReport R;
        uint64_t ev = 0;
        if (disc) R = disc->validate(path, reverse != 0, &st);
        else if (prog) R = prog->validate(path, reverse != 0, &st, &ev);
        else R = dich->validate(path, reverse != 0, &st, &ev);
        valid[i] = R.valid ? 1 : 0;
        if (t_fail) t_fail[i] = R.parameter;
        if (t_last_valid) t_last_valid[i] = R.lastValid;
        if (pair) pair[i] = R.index;
        if (reason) reason[i] = (uint8_t)R.reason;
        if (samples) samples[i] = R.samples;
        evals += ev; nsamples += R.samples;
      }
    } catch (std::exception& e) {
#pragma omp critical
      { g_err = e.what(); err = 2; }
    }
    calls += st.calls; iters += st.iterations; closed += st.closed;
  }
  h->pairEvals += evals; h->samples += nsamples;
  h->stats.calls += calls; h->stats.iterations += iters; h->stats.closed += closed;
  return err;
}

extern "C" {

int oracle_validate_edges(OracleHandle* h, int method, double param, int reverse, const double* q1, const double* q2,
                          const double* L, size_t n, uint32_t flags, uint8_t* valid, double* t_fail, double* t_last_valid,
                          int32_t* pair, uint8_t* reason, uint64_t* samples, int order_mode, uint64_t max_iter, int nthreads) {
  const Device& d = *h->dev;
  return validatePathsImpl(h, method, param, reverse, n, flags, valid, t_fail, t_last_valid, pair, reason, samples, order_mode, max_iter, nthreads,
                           [&](size_t i) { return StraightPath(d, q1 + i * d.nq, q2 + i * d.nq, L[i]); });
}

// Spline<BernsteinBasis, 3> edges (tests/test-continuous-validation.cc:288-427): base [n][nq], params [n][4][nv], L [n]
int oracle_validate_spline_edges(OracleHandle* h, int method, double param, int reverse, const double* base, const double* params,
                                 const double* L, size_t n, uint32_t flags, uint8_t* valid, double* t_fail, double* t_last_valid,
                                 int32_t* pair, uint8_t* reason, uint64_t* samples, uint64_t max_iter, int nthreads) {
  const Device& d = *h->dev;
  if (d.nq != d.nv) { g_err = "spline paths: vector-space configuration spaces only"; return 2; }
  return validatePathsImpl(h, method, param, reverse, n, flags, valid, t_fail, t_last_valid, pair, reason, samples, ORDER_RESET_PER_EDGE, max_iter,
                           nthreads, [&](size_t i) { return SplinePath(d, base + i * d.nq, params + i * 4 * (size_t)d.nv, L[i]); });
}
int oracle_spline_eval(OracleHandle* h, const double* base, const double* params, const double* L, const double* t, size_t n, double* q) {
  const Device& d = *h->dev;
  if (d.nq != d.nv) { g_err = "spline paths: vector-space configuration spaces only"; return 2; }
  for (size_t i = 0; i < n; ++i) SplinePath(d, base + i * d.nq, params + i * 4 * (size_t)d.nv, L[i]).eval(q + i * d.nq, t[i]);
  return 0;
}
// Path::velocityBound(t0, t1) of spline i: vb [n][nv]
int oracle_spline_velocity_bound(OracleHandle* h, const double* base, const double* params, const double* L, const double* ta, const double* tb,
                                 size_t n, double* vb) {
  const Device& d = *h->dev;
  if (d.nq != d.nv) { g_err = "spline paths: vector-space configuration spaces only"; return 2; }
  for (size_t i = 0; i < n; ++i) SplinePath(d, base + i * d.nq, params + i * 4 * (size_t)d.nv, L[i]).velocityBound(vb + i * d.nv, ta[i], tb[i]);
  return 0;
}

// ContinuousValidation::validate on a PathVector of StraightPaths (src/continuous-validation.cc:153-192), restated as the
// reference runs it: sub-path after sub-path, stop at the first invalid one, offset its report by the accumulated `param`.
// t_last_valid = end (reverse: start) of the valid part in the PathVector's parameterisation.
int oracle_validate_edge_chains(OracleHandle* h, int method, double param, int reverse, const double* q1, const double* q2,
                                const double* L, const int32_t* chain_begin, size_t n_chains, uint8_t* valid, double* t_fail,
                                double* t_last_valid, int32_t* pair, uint8_t* reason, int32_t* first_bad, int nthreads) {
  const Device& d = *h->dev;
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  int err = 0;
#pragma omp parallel num_threads(nthreads)
  {
    try {
      std::unique_ptr<Progressive> prog;
      std::unique_ptr<Dichotomy> dich;
      if (method == HPPB200_METHOD_PROGRESSIVE) prog.reset(new Progressive(d, param));
      else if (method == HPPB200_METHOD_DICHOTOMY) dich.reset(new Dichotomy(d, param));
      else throw std::runtime_error("PathVector recursion: continuous validations only");
      auto validateLeaf = [&](long i) {
        StraightPath path(d, q1 + (size_t)i * d.nq, q2 + (size_t)i * d.nq, L[i]);
        return prog ? prog->validate(path, reverse != 0) : dich->validate(path, reverse != 0);
      };
#pragma omp for schedule(dynamic, 4)
      for (long c = 0; c < (long)n_chains; ++c) {
        const long b = chain_begin[c], e = chain_begin[c + 1];
        bool ok = true;
        double tf = std::nan(""), tl = 0;
        int pr = -1, rs = HPPB200_REASON_VALID;
        long bad = -1;
        if (!reverse) {
          double prm = 0;  // :180
          for (long i = b; i < e; ++i) {
            PathReport R = validateLeaf(i);
            if (R.valid) prm += L[i];  // localPath->length()
            else { ok = false; tf = R.parameter + prm; tl = prm + R.lastValid; pr = R.index; rs = R.reason; bad = i; break; }
          }
          if (ok) tl = prm;
        } else {
          double prm = 0;  // path->length(): the PathVector's time range grows by each appended length
          for (long i = b; i < e; ++i) prm += L[i];
          for (long i = e; i != b; --i) {
            PathReport R = validateLeaf(i - 1);
            if (R.valid) prm -= L[i - 1];
            else { ok = false; tf = R.parameter + (prm - L[i - 1]); tl = (prm - L[i - 1]) + R.lastValid; pr = R.index; rs = R.reason; bad = i - 1; break; }
          }
          if (ok) tl = 0;
        }
        valid[c] = ok ? 1 : 0;
        if (t_fail) t_fail[c] = tf;
        if (t_last_valid) t_last_valid[c] = tl;
        if (pair) pair[c] = pr;
        if (reason) reason[c] = (uint8_t)rs;
        if (first_bad) first_bad[c] = (int32_t)bad;
      }
    } catch (std::exception& e) {
#pragma omp critical
      { g_err = e.what(); err = 2; }
    }
  }
  return err;
}

void oracle_get_counters(OracleHandle* h, uint64_t* out5) {
  out5[0] = h->stats.calls; out5[1] = h->stats.iterations; out5[2] = h->pairEvals; out5[3] = h->samples; out5[4] = h->stats.closed;
}

// ---- intervals (pinned against oracle/_ref) ----------------------------------------
// Applies n unions, writes the resulting list (max cap intervals) and returns its size.
int oracle_intervals_union(const double* lo, const double* hi, size_t n, double* out_lo, double* out_hi, int cap) {
  Intervals I;
  for (size_t i = 0; i < n; ++i) I.unionInterval(interval_t(lo[i], hi[i]));
  int k = 0;
  for (auto& s : I.list()) { if (k < cap) { out_lo[k] = s.first; out_hi[k] = s.second; } ++k; }
  return k;
}

// ---- nearest neighbour -----------------------------------------------------------
void oracle_knn_append(OracleHandle* h, const double* q, const int32_t* cc, size_t n) {
  for (size_t i = 0; i < n; ++i) h->roadmap->addNode(q + i * h->dev->nq, cc ? cc[i] : 0);
}
void oracle_knn_relabel(OracleHandle* h, int from, int to) {
  for (auto& c : h->roadmap->cc) if (c == from) c = to;
}
void oracle_knn_query(OracleHandle* h, const double* queries, size_t nqueries, int k, int cc, int64_t* idx, double* dist, int nthreads) {
  const Roadmap& rm = *h->roadmap;
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
  for (long i = 0; i < (long)nqueries; ++i) {
    double dk;
    const double* q = queries + (size_t)i * h->dev->nq;
    if (k == 1) {
      double d;
      long r = nnSearch(rm, q, cc, d);
      idx[i] = r; dist[i] = d;
    } else {
      auto res = nnKnearest(rm, q, cc, (size_t)k, dk);
      for (int j = 0; j < k; ++j) {
        idx[(size_t)i * k + j] = j < (int)res.size() ? res[j].second : -1;
        dist[(size_t)i * k + j] = j < (int)res.size() ? res[j].first : kInf;
      }
    }
  }
}
int64_t oracle_knn_within_ball(OracleHandle* h, const double* q, double radius, int cc, int64_t* idx, size_t cap) {
  auto r = nnWithinBall(*h->roadmap, q, cc, radius);
  for (size_t i = 0; i < r.size() && i < cap; ++i) idx[i] = r[i];
  return (int64_t)r.size();
}

}  // extern "C"
