// Model compiler: flattens a robot + obstacle description into the device tables, in the
// exact pair ORDER hpp-core would use.  Host-side restatement of
//   ObstacleUser::addRobotCollisionPairs / addObstacle / setSecurityMargins
//       (reference src/obstacle-user.cc:222-235, :53-76, :166-196)
//   ContinuousValidation::Initialize::doExecute / AddObstacle::doExecute
//       (reference src/continuous-validation.cc:63-98, :103-118)
//   SolidSolidCollision::Model::{computeSequenceOfJoints,computeCoefficients,setCoefficients}
//       (reference src/continuous-validation/solid-solid-collision.cc:162-244)
//   hpp-pinocchio Body::radius / Joint::{maximalDistanceToParent,upperBound*Velocity}
//       
// Setup-time integer/FP64 work; runs once per model on the host.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <functional>

#include "engine.hpp"

namespace hppb200 {

static int pairKind(int ta, int tb) {  // ta <= tb
  if (ta == HPPB200_SHAPE_SPHERE && tb == HPPB200_SHAPE_SPHERE) return 0;
  if (ta == HPPB200_SHAPE_SPHERE && tb == HPPB200_SHAPE_CAPSULE) return 1;
  if (ta == HPPB200_SHAPE_CAPSULE && tb == HPPB200_SHAPE_CAPSULE) return 2;
  if (ta == HPPB200_SHAPE_SPHERE && tb == HPPB200_SHAPE_BOX) return 3;
  return 4;
}

static DevShapeH makeShape(int type, const double* p, int vb, int vc) {
  DevShapeH s{};
  if (type < HPPB200_SHAPE_SPHERE || type > HPPB200_SHAPE_CONVEX) throw std::invalid_argument("unknown shape type");
  s.type = type;
  for (int k = 0; k < 4; ++k) s.p[k] = p[k];
  s.vbegin = type == HPPB200_SHAPE_CONVEX ? vb : 0;
  s.vcount = type == HPPB200_SHAPE_CONVEX ? vc : 0;
  return s;
}

static const double kCullSlack = 1e-9;  // >> FP64 rounding of the bound, << anything physical

static DevBound makeBound(const DevShapeH& s, const std::vector<double>& verts) {
  DevBound b{};
  const double up = 1.0 + 1e-12;  // round radii up
  if (s.type == HPPB200_SHAPE_SPHERE || s.type == HPPB200_SHAPE_CAPSULE) b.br = s.p[0];
  else if (s.type == HPPB200_SHAPE_BOX) b.br = std::sqrt(s.p[0] * s.p[0] + s.p[1] * s.p[1] + s.p[2] * s.p[2]) * up;
  else {
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = 0; i < s.vcount; ++i)
      for (int k = 0; k < 3; ++k) {
        double c = verts[3 * (size_t)(s.vbegin + i) + k];
        lo[k] = std::min(lo[k], c);
        hi[k] = std::max(hi[k], c);
      }
    for (int k = 0; k < 3; ++k) b.bc[k] = s.vcount ? 0.5 * (lo[k] + hi[k]) : 0.0;
    double r2 = 0;
    for (int i = 0; i < s.vcount; ++i) {
      double d2 = 0;
      for (int k = 0; k < 3; ++k) { double d = verts[3 * (size_t)(s.vbegin + i) + k] - b.bc[k]; d2 += d * d; }
      r2 = std::max(r2, d2);
    }
    b.br = std::sqrt(r2) * up;
  }
  return b;
}

static double norm3(double x, double y, double z) { return std::sqrt(std::fma(z, z, std::fma(y, y, x * x))); }
// |R p + t| with the engine's dot convention
static double actNorm(const double* P, double x, double y, double z) {
  double w[3];
  for (int i = 0; i < 3; ++i) w[i] = std::fma(P[3 * i + 2], z, std::fma(P[3 * i + 1], y, P[3 * i] * x)) + P[9 + i];
  return norm3(w[0], w[1], w[2]);
}

Model::Model(const hppb200_model_desc& d) {
  if (d.njoints <= 0 || !d.joint_parent || !d.joint_type || !d.joint_placement || !d.q_lower || !d.q_upper)
    throw std::invalid_argument("model description: missing joint arrays");
  njoints = d.njoints;
  joints.resize(njoints);
  joint_nq.resize(njoints);
  joint_nv.resize(njoints);
  for (int k = 0; k < njoints; ++k) {
    DevJoint& J = joints[k];
    J.parent = d.joint_parent[k];
    if (J.parent < 0 || J.parent > k) throw std::invalid_argument("joint parent must precede its child");
    J.type = d.joint_type[k];
    if (J.type < HPPB200_JOINT_REVOLUTE_X || J.type > HPPB200_JOINT_FREEFLYER) throw std::invalid_argument("unknown joint type");
    std::memcpy(J.P, d.joint_placement + 12 * k, sizeof(J.P));
    joint_nq[k] = J.type == HPPB200_JOINT_FREEFLYER ? 7 : 1;
    joint_nv[k] = J.type == HPPB200_JOINT_FREEFLYER ? 6 : 1;
    J.idx_q = nq;
    J.idx_v = nv;
    nq += joint_nq[k];
    nv += joint_nv[k];
  }
  lower.assign(d.q_lower, d.q_lower + nq);
  upper.assign(d.q_upper, d.q_upper + nq);
  if (d.nvertices > 0 && d.vertices) vertices.assign(d.vertices, d.vertices + 3 * (size_t)d.nvertices);
  for (int g = 0; g < d.ngeoms; ++g) {
    DevGeom G{};
    G.joint = d.geom_joint[g];
    if (G.joint < 0 || G.joint > njoints) throw std::invalid_argument("geometry parent joint out of range");
    std::memcpy(G.P, d.geom_placement + 12 * g, sizeof(G.P));
    G.shape = makeShape(d.geom_shape[g], d.geom_params + 4 * g, d.geom_vert_begin ? d.geom_vert_begin[g] : 0,
                        d.geom_vert_count ? d.geom_vert_count[g] : 0);
    G.bound = makeBound(G.shape, vertices);
    geoms.push_back(G);
  }
  for (int o = 0; o < d.nobstacles; ++o) {
    DevObstacle O{};
    std::memcpy(O.P, d.obst_placement + 12 * o, sizeof(O.P));
    O.shape = makeShape(d.obst_shape[o], d.obst_params + 4 * o, d.obst_vert_begin ? d.obst_vert_begin[o] : 0,
                        d.obst_vert_count ? d.obst_vert_count[o] : 0);
    O.bound = makeBound(O.shape, vertices);
    for (int i = 0; i < 3; ++i)
      O.wc[i] = std::fma(O.P[3 * i + 2], O.bound.bc[2], std::fma(O.P[3 * i + 1], O.bound.bc[1], O.P[3 * i] * O.bound.bc[0])) + O.P[9 + i];
    obstacles.push_back(O);
  }
  if (d.security_margins) {
    size_t n = (size_t)(njoints + 1);
    margins.assign(d.security_margins, d.security_margins + n * n);
  }
  weights.assign(njoints, 1.0);
  if (d.weights) weights.assign(d.weights, d.weights + njoints);
  break_distance = d.break_distance;

  // Body::radius per joint (pinocchio computeBodyRadius): farthest corner of each geometry's
  // local AABB mapped by its placement.
  radii.assign(njoints + 1, 0.0);
  for (const DevGeom& G : geoms) {
    double lo[3], hi[3];
    const DevShapeH& s = G.shape;
    if (s.type == HPPB200_SHAPE_SPHERE) { for (int k = 0; k < 3; ++k) { lo[k] = -s.p[0]; hi[k] = s.p[0]; } }
    else if (s.type == HPPB200_SHAPE_CAPSULE) { lo[0] = lo[1] = -s.p[0]; hi[0] = hi[1] = s.p[0]; lo[2] = -(s.p[1] + s.p[0]); hi[2] = s.p[1] + s.p[0]; }
    else if (s.type == HPPB200_SHAPE_BOX) { for (int k = 0; k < 3; ++k) { lo[k] = -s.p[k]; hi[k] = s.p[k]; } }
    else {
      for (int k = 0; k < 3; ++k) { lo[k] = INFINITY; hi[k] = -INFINITY; }
      for (int i = 0; i < s.vcount; ++i)
        for (int k = 0; k < 3; ++k) {
          double c = vertices[3 * (size_t)(s.vbegin + i) + k];
          lo[k] = std::min(lo[k], c);
          hi[k] = std::max(hi[k], c);
        }
    }
    for (int c = 0; c < 8; ++c) {
      double n = actNorm(G.P, (c & 1) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1], (c & 4) ? hi[2] : lo[2]);
      if (n > radii[G.joint]) radii[G.joint] = n;
    }
  }
  buildConfigPairs(d);
  buildBodyPairs(d);
  applyPairOps(d);
  if (d.check_parameterized) {  // CollisionValidation::validate tests pPairs_ after cPairs_ (collision-validation.cc:62-67)
    nParameterized = (int)pPairs.size();
    cPairs.insert(cPairs.end(), pPairs.begin(), pPairs.end());
  }
  buildBlob();
}

// The ObstacleUserInterface calls an hpp-core validator received after construction, replayed in order on the two
// tables (configuration pairs: reference src/obstacle-user.cc:60-220; continuous body pairs:
// src/continuous-validation.cc:224-367, solid-solid-collision.cc:101-122).
void Model::applyPairOps(const hppb200_model_desc& d) {
  if (d.npair_ops <= 0) return;
  if (!d.pair_ops) throw std::invalid_argument("pair_ops is null");
  const int n1 = njoints + 1;
  auto jointOf = [&](int ref) { return ref >= 0 ? geoms[ref].joint : 0; };
  auto checkObstacle = [&](int o) { if (o < 0 || o >= (int)obstacles.size()) throw std::invalid_argument("pair op: obstacle index out of range"); };
  auto checkJoint = [&](int j) { if (j < 1 || j > njoints) throw std::invalid_argument("pair op: joint index out of range"); };
  for (int k = 0; k < d.npair_ops; ++k) {
    const hppb200_pair_op& op = d.pair_ops[k];
    switch (op.kind) {
      case HPPB200_OP_ADD_OBSTACLE_TO_JOINT: {  // obstacle-user.cc:60-76
        checkObstacle(op.a); checkJoint(op.b);
        std::vector<int> stack{op.b};
        // recursion of the reference: the joint's inner objects, then each child joint in order (depth first)
        std::function<void(int)> add = [&](int j) {
          for (int g = 0; g < (int)geoms.size(); ++g)
            if (geoms[g].joint == j) cPairs.push_back(DevPair{g, -(op.a + 1), 0, 0, 0.0, 0.0});
          if (op.flag)
            for (int c = 1; c <= njoints; ++c)
              if (joints[c - 1].parent == j) add(c);
        };
        add(op.b);
        break;
      }
      case HPPB200_OP_ADD_OBSTACLE: {  // obstacle-user.cc:53-58 (default request: margin 0), continuous-validation.cc:103-118
        checkObstacle(op.a);
        for (int j = 1; j <= njoints; ++j) {
          BodyPairHost bp;
          bp.joint_a = j;
          bp.joint_b = 0;
          for (int g = 0; g < (int)geoms.size(); ++g)
            if (geoms[g].joint == j) {
              cPairs.push_back(DevPair{g, -(op.a + 1), 0, 0, 0.0, 0.0});
              bp.pairs.push_back(DevPair{g, -(op.a + 1), 0, 0, 0.0, 0.0});
            }
          if (bp.pairs.empty()) continue;
          finishBodyPair(bp);
          bodyPairs.push_back(bp);
        }
        break;
      }
      case HPPB200_OP_REMOVE_OBSTACLE_FROM_JOINT: {  // obstacle-user.cc:87-118, continuous-validation.cc:224-252
        checkObstacle(op.a); checkJoint(op.b);
        const int oref = -(op.a + 1);
        for (int g = 0; g < (int)geoms.size(); ++g) {
          if (geoms[g].joint != op.b) continue;
          size_t deleted = 0;
          for (size_t i = 0; i < cPairs.size();)
            if (cPairs[i].a == g && cPairs[i].b == oref) { cPairs.erase(cPairs.begin() + i); ++deleted; }
            else ++i;
          if (!deleted) throw std::runtime_error("removeObstacleFromJoint: the obstacle is not registered as obstacle for the joint");
        }
        bool removed = false, hasBody = false;
        for (const DevGeom& G : geoms) hasBody = hasBody || G.joint == op.b;
        for (size_t i = 0; i < bodyPairs.size();) {
          BodyPairHost& bp = bodyPairs[i];
          bool erased = false;
          if (bp.joint_a == op.b && bp.joint_b == 0) {
            const size_t before = bp.pairs.size();
            bp.pairs.erase(std::remove_if(bp.pairs.begin(), bp.pairs.end(), [&](const DevPair& p) { return p.b == oref; }), bp.pairs.end());
            if (bp.pairs.size() != before) {
              removed = true;
              if (bp.pairs.empty()) { bodyPairs.erase(bodyPairs.begin() + i); erased = true; }
            }
          }
          if (!erased) ++i;
        }
        if (hasBody && !removed) throw std::runtime_error("removeObstacleFromJoint: the obstacle is not registered as obstacle for the joint");
        break;
      }
      case HPPB200_OP_FILTER_COLLISION_PAIRS: {  // obstacle-user.cc:120-164, continuous-validation.cc:266-300
        if (!op.data) throw std::invalid_argument("filterCollisionPairs: null relative-motion matrix");
        const int32_t* M = (const int32_t*)op.data;
        for (size_t i = 0; i < cPairs.size();) {
          const int m = M[(size_t)jointOf(cPairs[i].a) * n1 + jointOf(cPairs[i].b)];
          if (m == 1) { pPairs.push_back(cPairs[i]); cPairs.erase(cPairs.begin() + i); }
          else if (m == 0) { dPairs.push_back(cPairs[i]); cPairs.erase(cPairs.begin() + i); }
          else ++i;
        }
        for (size_t i = 0; i < bodyPairs.size();) {
          const int m = M[(size_t)bodyPairs[i].joint_a * n1 + bodyPairs[i].joint_b];
          if (m == 0 || m == 1) bodyPairs.erase(bodyPairs.begin() + i);  // "Parameterized collision pairs treated as Constrained"
          else ++i;
        }
        break;
      }
      case HPPB200_OP_SET_SECURITY_MARGINS: {  // obstacle-user.cc:166-196, continuous-validation.cc:302-330
        if (!op.data) throw std::invalid_argument("setSecurityMargins: null matrix");
        const double* M = (const double*)op.data;
        for (std::vector<DevPair>* tab : {&cPairs, &pPairs, &dPairs})
          for (DevPair& p : *tab) p.margin = M[(size_t)jointOf(p.a) * n1 + jointOf(p.b)];
        for (BodyPairHost& bp : bodyPairs)
          for (DevPair& p : bp.pairs) p.margin = M[(size_t)bp.joint_a * n1 + bp.joint_b];
        break;
      }
      case HPPB200_OP_SET_MARGIN_BETWEEN_BODIES: {  // obstacle-user.cc:198-220, continuous-validation.cc:332-367
        auto matches = [&](const DevPair& p) { return (p.a == op.a && p.b == op.b) || (p.a == op.b && p.b == op.a); };
        bool found = false;
        for (std::vector<DevPair>* tab : {&cPairs, &pPairs, &dPairs}) {
          for (DevPair& p : *tab)
            if (matches(p)) { p.margin = op.value; found = true; break; }  // first match only
          if (found) break;
        }
        for (BodyPairHost& bp : bodyPairs)  // every match; the reference never reports a miss here (found starts true)
          for (DevPair& p : bp.pairs)
            if (matches(p)) p.margin = op.value;
        if (!found) throw std::invalid_argument("setSecurityMarginBetweenBodies: could not find a collision pair between the two bodies");
        break;
      }
      default: throw std::invalid_argument("unknown pair op");
    }
  }
}

// cPairs_ order: active self pairs, then for every obstacle (call order) every moving joint
// (index order) every inner object of its body.  Geometry on the universe gets no obstacle pair.
void Model::buildConfigPairs(const hppb200_model_desc& d) {
  for (int p = 0; p < d.nself_pairs; ++p) {
    int a = d.self_pairs[2 * p], b = d.self_pairs[2 * p + 1];
    if (a < 0 || b < 0 || a >= (int)geoms.size() || b >= (int)geoms.size()) throw std::invalid_argument("self pair index out of range");
    cPairs.push_back(DevPair{a, b, 0, 0, margin(geoms[a].joint, geoms[b].joint), 0.0});
  }
  for (int o = 0; o < (int)obstacles.size(); ++o) {
    if (d.obst_global && !d.obst_global[o]) continue;  // only paired through HPPB200_OP_ADD_OBSTACLE_TO_JOINT
    for (int j = 1; j <= njoints; ++j)
      for (int g = 0; g < (int)geoms.size(); ++g)
        if (geoms[g].joint == j) cPairs.push_back(DevPair{g, -(o + 1), 0, 0, margin(j, 0), 0.0});
  }
}

// Joint sequence and velocity coefficients of one body pair (SolidSolidCollision::Model::computeSequenceOfJoints /
// computeCoefficients / setCoefficients, reference src/continuous-validation/solid-solid-collision.cc:162-244)
void Model::finishBodyPair(BodyPairHost& bp) const {
  auto parentOf = [&](int j) { return joints[j - 1].parent; };
  auto linVel = [&](int j) { int t = joints[j - 1].type; return (t >= HPPB200_JOINT_PRISMATIC_X) ? 1.0 : 0.0; };
  auto angVel = [&](int j) { int t = joints[j - 1].type; return (t <= HPPB200_JOINT_REVOLUTE_Z || t == HPPB200_JOINT_FREEFLYER) ? 1.0 : 0.0; };
  auto maxDistToParent = [&](int j) {
    const DevJoint& J = joints[j - 1];
    if (J.type <= HPPB200_JOINT_REVOLUTE_Z) return norm3(J.P[9], J.P[10], J.P[11]);
    if (J.type == HPPB200_JOINT_FREEFLYER) {  // farthest corner of the translation bounds (inf if unbounded)
      double best = 0;
      for (int c = 0; c < 8; ++c) {
        const double p[3] = {(c & 1) ? upper[J.idx_q] : lower[J.idx_q], (c & 2) ? upper[J.idx_q + 1] : lower[J.idx_q + 1],
                             (c & 4) ? upper[J.idx_q + 2] : lower[J.idx_q + 2]};
        auto row = [&](int i) { return std::fma(J.P[3 * i + 2], p[2], std::fma(J.P[3 * i + 1], p[1], J.P[3 * i] * p[0])); };
        const double n = norm3(row(0) + J.P[9], row(1) + J.P[10], row(2) + J.P[11]);
        if (!(n <= best)) best = n;
      }
      return best;
    }
    int a = J.type - HPPB200_JOINT_PRISMATIC_X;  // prismatic: farthest end of the stroke
    double best = 0;
    for (double s : {lower[J.idx_q], upper[J.idx_q]}) {
      double n = norm3(std::fma(s, J.P[a], J.P[9]), std::fma(s, J.P[3 + a], J.P[10]), std::fma(s, J.P[6 + a], J.P[11]));
      best = std::max(best, n);
    }
    return best;
  };
  auto coefficients = [&](const std::vector<int>& seq, int joint_a) {
    std::vector<std::pair<int, double>> out;
    double cumulative = radii[joint_a];  // radius of body a for both directions, as the reference does
    for (size_t i = 0; i + 1 < seq.size(); ++i) {
      int child;
      if (seq[i] != 0 && parentOf(seq[i]) == seq[i + 1]) child = seq[i];
      else if (seq[i + 1] != 0 && parentOf(seq[i + 1]) == seq[i]) child = seq[i + 1];
      else throw std::runtime_error("joint sequence is not a kinematic chain");
      out.push_back({child, linVel(child) + cumulative * angVel(child)});
      cumulative += maxDistToParent(child);
    }
    return out;
  };
    std::vector<int> up, down;  // climb from the larger index until both meet
    int ia = bp.joint_a, ib = bp.joint_b;
    while (ia != ib) {
      if (ia > ib) { up.push_back(ia); ia = parentOf(ia); }
      else { down.push_back(ib); ib = parentOf(ib); }
    }
    up.push_back(ia);
    std::vector<int> seq(up);
    seq.insert(seq.end(), down.rbegin(), down.rend());
    bp.coefs = coefficients(seq, bp.joint_a);
    if (bp.joint_b != 0) {
      std::reverse(seq.begin(), seq.end());
      bp.rcoefs = coefficients(seq, bp.joint_a);
    }
}

void Model::buildBodyPairs(const hppb200_model_desc& d) {
  auto finish = [&](BodyPairHost& bp) { finishBodyPair(bp); };
  // self pairs: one body pair per ordered joint pair, in order of first appearance
  std::map<std::pair<int, int>, size_t> seen;
  for (int p = 0; p < d.nself_pairs; ++p) {
    int ga = d.self_pairs[2 * p], gb = d.self_pairs[2 * p + 1];
    std::pair<int, int> jp(geoms[ga].joint, geoms[gb].joint);
    if (jp.first == jp.second) continue;
    auto it = seen.find(jp);
    if (it == seen.end()) {
      BodyPairHost bp;
      int j1 = jp.first, j2 = jp.second;
      if (j2 == 0) std::swap(j1, j2);
      bp.joint_a = j2;
      bp.joint_b = j1;
      finish(bp);
      bodyPairs.push_back(bp);
      it = seen.insert({jp, bodyPairs.size() - 1}).first;
    }
    bodyPairs[it->second].pairs.push_back(DevPair{ga, gb, 0, 0, margin(jp.first, jp.second), 0.0});
  }
  // obstacles: one body pair per (joint carrying a body, obstacle)
  for (int o = 0; o < (int)obstacles.size(); ++o)
    for (int j = 1; j <= njoints; ++j) {
      if (d.obst_global && !d.obst_global[o]) continue;
      BodyPairHost bp;
      bp.joint_a = j;
      bp.joint_b = 0;
      for (int g = 0; g < (int)geoms.size(); ++g)
        if (geoms[g].joint == j) bp.pairs.push_back(DevPair{g, -(o + 1), 0, 0, margin(j, 0), 0.0});
      if (bp.pairs.empty()) continue;
      finish(bp);
      bodyPairs.push_back(bp);
    }
}

void Model::buildBlob() {
  auto shapeOf = [&](int ref) -> const DevShapeH& { return ref >= 0 ? geoms[ref].shape : obstacles[-(ref + 1)].shape; };
  auto canonical = [&](DevPair p) {
    if (shapeOf(p.a).type > shapeOf(p.b).type) std::swap(p.a, p.b);
    p.kind = pairKind(shapeOf(p.a).type, shapeOf(p.b).type);
    // conservative cull: cores farther apart than the inflation radii + margin (+ slack) cannot collide
    p.bound = shapeOf(p.a).type == HPPB200_SHAPE_CAPSULE ? 1 : 0;
    const double ra = p.a >= 0 ? geoms[p.a].bound.br : obstacles[-(p.a + 1)].bound.br;
    const double rb = p.b >= 0 ? geoms[p.b].bound.br : obstacles[-(p.b + 1)].bound.br;
    const double reach = std::max(0.0, ra + rb + p.margin + kCullSlack);
    p.cull2 = reach * reach * (1.0 + 1e-12);
    return p;
  };
  auto append = [&](const void* src, size_t bytes) {
    size_t off = blob.size();
    blob.resize(off + ((bytes + 15) & ~size_t(15)), 0);
    if (bytes) std::memcpy(blob.data() + off, src, bytes);
    return (uint32_t)off;
  };
  std::vector<DevPair> cp, bg;
  std::vector<DevBodyPair> bps;
  std::vector<DevCoef> cf, rcf;
  for (const DevPair& p : cPairs) cp.push_back(canonical(p));
  // DistanceBetweenObjects::addObstacle (src/distance-between-objects.cc:41-53): obstacle-major, joints in order, inner objects
  dboPairs.clear();
  for (int o = 0; o < (int)obstacles.size(); ++o)
    for (int j = 1; j <= njoints; ++j)
      for (int g = 0; g < (int)geoms.size(); ++g)
        if (geoms[g].joint == j) dboPairs.push_back(canonical(DevPair{g, -(o + 1), 0, 0, 0.0, 0.0}));
  for (const BodyPairHost& b : bodyPairs) {
    DevBodyPair B{};
    B.joint_a = b.joint_a;
    B.joint_b = b.joint_b;
    B.pair_begin = (int)bg.size();
    for (const DevPair& p : b.pairs) {
      bg.push_back(canonical(p));
      if (bg.back().kind == 4 /* PK_GJK */) B.has_gjk = 1;
    }
    B.pair_end = (int)bg.size();
    B.coef_begin = (int)cf.size();
    for (auto& c : b.coefs) cf.push_back(DevCoef{joints[c.first - 1].idx_v, (int16_t)joint_nv[c.first - 1], (int16_t)joints[c.first - 1].idx_q, c.second});
    B.coef_end = (int)cf.size();
    B.rcoef_begin = (int)rcf.size();
    for (auto& c : b.rcoefs) rcf.push_back(DevCoef{joints[c.first - 1].idx_v, (int16_t)joint_nv[c.first - 1], (int16_t)joints[c.first - 1].idx_q, c.second});
    B.rcoef_end = (int)rcf.size();
    bps.push_back(B);
  }
  view = ModelView{};
  view.off_joints = append(joints.data(), joints.size() * sizeof(DevJoint));
  view.off_geoms = append(geoms.data(), geoms.size() * sizeof(DevGeom));
  view.off_obstacles = append(obstacles.data(), obstacles.size() * sizeof(DevObstacle));
  view.off_bodypairs = append(bps.data(), bps.size() * sizeof(DevBodyPair));
  view.off_coefs = append(cf.data(), cf.size() * sizeof(DevCoef));
  view.off_rcoefs = append(rcf.data(), rcf.size() * sizeof(DevCoef));
  // (the body pairs' geometry pairs sit right in front of the bounds: the first-sample screen of the edge path stages
  //  [0, off_bodypairs) + [off_bgpairs, off_cpairs) and reads them as ITS pair table, kernels_config.cu::configPassSetup)
  view.off_bgpairs = append(bg.data(), bg.size() * sizeof(DevPair));
  view.off_lower = append(lower.data(), lower.size() * sizeof(double));
  view.off_upper = append(upper.data(), upper.size() * sizeof(double));
  view.off_weights = append(weights.data(), weights.size() * sizeof(double));
  // the configuration pair table comes last: the edge engine stages only the prefix before it
  view.edge_blob_bytes = (uint32_t)blob.size();
  view.off_cpairs = append(cp.data(), cp.size() * sizeof(DevPair));
  view.blob_bytes = (uint32_t)blob.size();
  view.njoints = njoints;
  view.ngeoms = (int)geoms.size();
  view.nobstacles = (int)obstacles.size();
  view.ncpairs = (int)cp.size();
  view.nbodypairs = (int)bps.size();
  view.nbgpairs = (int)bg.size();
  view.ncoefs = (int)cf.size();
  view.nrcoefs = (int)rcf.size();
  view.nq = nq;
  view.nv = nv;
}

ModelView Model::deviceView(int device) const {
  std::lock_guard<std::mutex> lock(mu_);
  auto it = dev_.find(device);
  if (it == dev_.end()) {
    HPPB200_CUDA_CHECK(cudaSetDevice(device));
    void* dblob = nullptr;
    void* dverts = nullptr;
    HPPB200_CUDA_CHECK(cudaMalloc(&dblob, blob.size()));
    HPPB200_CUDA_CHECK(cudaMemcpy(dblob, blob.data(), blob.size(), cudaMemcpyHostToDevice));
    if (!vertices.empty()) {
      HPPB200_CUDA_CHECK(cudaMalloc(&dverts, vertices.size() * sizeof(double)));
      HPPB200_CUDA_CHECK(cudaMemcpy(dverts, vertices.data(), vertices.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    it = dev_.insert({device, {dblob, dverts}}).first;
  }
  ModelView v = view;
  v.blob = (const unsigned char*)it->second.first;
  v.vertices = (const double*)it->second.second;
  return v;
}

Model::~Model() {
  for (auto& kv : dev_) {
    cudaSetDevice(kv.first);
    if (kv.second.first) cudaFree(kv.second.first);
    if (kv.second.second) cudaFree(kv.second.second);
  }
}

}  // namespace hppb200
