// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_math.hpp header).
//
// Robot / environment tables and forward kinematics as hpp-core sees them through
// hpp-pinocchio (Device, Joint, Body, CollisionObject) and pinocchio
// (forwardKinematics, updateGeometryPlacements).  PARITY UNPINNED: those libraries are
// absent from /root/reference; semantics per SURVEY.md Appendix A.2/A.3.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/hppb200.h"  // data format of the model description only
#include "oracle_narrowphase.hpp"

namespace oracle {

struct Joint {
  int parent = 0;  // pinocchio joint index of the parent (0 = universe)
  int type = 0;    // HPPB200_JOINT_*
  SE3 placement;   // jointPlacements[i]
  int idx_q = 0, idx_v = 0, nq = 1, nv = 1;
};

struct Geometry {
  int joint = 0;  // parent joint (0 = universe)
  SE3 placement;  // in the joint frame (obstacles: world placement)
  Shape shape;
};

// pinocchio::DeviceData equivalent: result of FK for one configuration
struct DeviceData {
  std::vector<SE3> oMi;  // [njoints+1], oMi[0] = identity
  std::vector<SE3> oMg;  // [ngeoms]
};

inline int jointNq(int type) { return type == HPPB200_JOINT_FREEFLYER ? 7 : 1; }
inline int jointNv(int type) { return type == HPPB200_JOINT_FREEFLYER ? 6 : 1; }

// One recorded ObstacleUserInterface call (include/hpp/core/obstacle-user.hh:47-88), see hppb200_pair_op
struct PairOp {
  int kind = 0, a = 0, b = 0, flag = 0;
  double value = 0;
  std::vector<int> relativeMotion;  // (njoints+1)^2, RelativeMotion::matrix_type
  std::vector<double> matrix;       // (njoints+1)^2 security margins
};

struct Device {
  std::vector<PairOp> pairOps;
  std::vector<unsigned char> obstGlobal;  // empty: every obstacle went through addObstacle()
  bool checkParameterized = false;
  bool isGlobal(size_t o) const { return obstGlobal.empty() || obstGlobal[o] != 0; }
  std::vector<Joint> joints;  // index 0 = universe placeholder
  std::vector<Geometry> geoms;
  std::vector<Geometry> obstacles;
  std::vector<std::pair<int, int>> selfPairs;  // active GeomModel::collisionPairs
  std::vector<double> lower, upper;            // configuration bounds [nq]
  std::vector<double> margins;                 // (njoints+1)^2 or empty
  std::vector<double> weights;                 // [njoints]
  std::vector<double> vertexPool;
  double breakDistance = 1e-2;
  int nq = 0, nv = 0;

  int njoints() const { return (int)joints.size() - 1; }
  double margin(int j1, int j2) const {
    return margins.empty() ? 0.0 : margins[(size_t)j1 * joints.size() + j2];
  }

  static Shape makeShape(int type, const double* p, const double* pool, int vb, int vc) {
    Shape s;
    s.type = type;
    for (int k = 0; k < 4; ++k) s.p[k] = p[k];
    if (type == CONVEX) { s.verts = pool + 3 * (size_t)vb; s.nverts = vc; }
    return s;
  }

  explicit Device(const hppb200_model_desc& d) {
    joints.resize(d.njoints + 1);
    joints[0].placement = SE3::identity();
    for (int i = 1; i <= d.njoints; ++i) {
      Joint& j = joints[i];
      j.parent = d.joint_parent[i - 1];
      if (j.parent < 0 || j.parent >= i) throw std::invalid_argument("joint parent must precede child");
      j.type = d.joint_type[i - 1];
      j.placement = SE3::from12(d.joint_placement + 12 * (i - 1));
      j.nq = jointNq(j.type); j.nv = jointNv(j.type);
      j.idx_q = nq; j.idx_v = nv;
      nq += j.nq; nv += j.nv;
    }
    lower.assign(d.q_lower, d.q_lower + nq);
    upper.assign(d.q_upper, d.q_upper + nq);
    if (d.nvertices > 0) vertexPool.assign(d.vertices, d.vertices + 3 * (size_t)d.nvertices);
    for (int g = 0; g < d.ngeoms; ++g) {
      Geometry G;
      G.joint = d.geom_joint[g];
      G.placement = SE3::from12(d.geom_placement + 12 * g);
      G.shape = makeShape(d.geom_shape[g], d.geom_params + 4 * g, vertexPool.data(),
                          d.geom_vert_begin ? d.geom_vert_begin[g] : 0, d.geom_vert_count ? d.geom_vert_count[g] : 0);
      geoms.push_back(G);
    }
    for (int o = 0; o < d.nobstacles; ++o) {
      Geometry G;
      G.joint = 0;
      G.placement = SE3::from12(d.obst_placement + 12 * o);
      G.shape = makeShape(d.obst_shape[o], d.obst_params + 4 * o, vertexPool.data(),
                          d.obst_vert_begin ? d.obst_vert_begin[o] : 0, d.obst_vert_count ? d.obst_vert_count[o] : 0);
      obstacles.push_back(G);
    }
    for (int p = 0; p < d.nself_pairs; ++p) selfPairs.push_back({d.self_pairs[2 * p], d.self_pairs[2 * p + 1]});
    if (d.security_margins) {
      size_t n = (size_t)(d.njoints + 1);
      margins.assign(d.security_margins, d.security_margins + n * n);
    }
    weights.assign(d.njoints, 1.0);
    if (d.weights) weights.assign(d.weights, d.weights + d.njoints);
    breakDistance = d.break_distance;
    if (d.obst_global) obstGlobal.assign(d.obst_global, d.obst_global + d.nobstacles);
    checkParameterized = d.check_parameterized != 0;
    const size_t nn = (size_t)(d.njoints + 1) * (size_t)(d.njoints + 1);
    for (int k = 0; k < d.npair_ops; ++k) {
      const hppb200_pair_op& o = d.pair_ops[k];
      PairOp op;
      op.kind = o.kind; op.a = o.a; op.b = o.b; op.flag = o.flag; op.value = o.value;
      if (o.kind == HPPB200_OP_FILTER_COLLISION_PAIRS && o.data) op.relativeMotion.assign((const int32_t*)o.data, (const int32_t*)o.data + nn);
      if (o.kind == HPPB200_OP_SET_SECURITY_MARGINS && o.data) op.matrix.assign((const double*)o.data, (const double*)o.data + nn);
      pairOps.push_back(op);
    }
  }

  // pinocchio joint transform M_joint(q)
  static SE3 jointTransform(const Joint& j, const double* q) {
    SE3 M = SE3::identity();
    switch (j.type) {
      case HPPB200_JOINT_REVOLUTE_X: case HPPB200_JOINT_REVOLUTE_Y: case HPPB200_JOINT_REVOLUTE_Z: {
        double s, c;
        sincos_portable(q[0], &s, &c);
        int a = j.type - HPPB200_JOINT_REVOLUTE_X, b = (a + 1) % 3, cc = (a + 2) % 3;
        M.R.m[b][b] = c; M.R.m[b][cc] = -s; M.R.m[cc][b] = s; M.R.m[cc][cc] = c;
        break;
      }
      case HPPB200_JOINT_PRISMATIC_X: M.t.x = q[0]; break;
      case HPPB200_JOINT_PRISMATIC_Y: M.t.y = q[0]; break;
      case HPPB200_JOINT_PRISMATIC_Z: M.t.z = q[0]; break;
      case HPPB200_JOINT_FREEFLYER: {
        M.t = {q[0], q[1], q[2]};
        double x = q[3], y = q[4], z = q[5], w = q[6];
        // Eigen::Quaternion::toRotationMatrix operation order
        double tx = 2 * x, ty = 2 * y, tz = 2 * z;
        double twx = tx * w, twy = ty * w, twz = tz * w;
        double txx = tx * x, txy = ty * x, txz = tz * x;
        double tyy = ty * y, tyz = tz * y, tzz = tz * z;
        M.R.m[0][0] = 1 - (tyy + tzz); M.R.m[0][1] = txy - twz; M.R.m[0][2] = txz + twy;
        M.R.m[1][0] = txy + twz; M.R.m[1][1] = 1 - (txx + tzz); M.R.m[1][2] = tyz - twx;
        M.R.m[2][0] = txz - twy; M.R.m[2][1] = tyz + twx; M.R.m[2][2] = 1 - (txx + tyy);
        break;
      }
      default: throw std::invalid_argument("unknown joint type");
    }
    return M;
  }

  // pinocchio::forwardKinematics + updateGeometryPlacements (Appendix A.2):
  //   liMi = placement * M_joint(q); oMi = oMi[parent] * liMi; oMg = oMi[joint] * placement_g
  void forwardKinematics(const double* q, DeviceData& data) const {
    data.oMi.resize(joints.size());
    data.oMg.resize(geoms.size());
    data.oMi[0] = SE3::identity();
    for (size_t i = 1; i < joints.size(); ++i) {
      const Joint& j = joints[i];
      SE3 liMi = compose(j.placement, jointTransform(j, q + j.idx_q));
      data.oMi[i] = j.parent == 0 ? liMi : compose(data.oMi[j.parent], liMi);
    }
    for (size_t g = 0; g < geoms.size(); ++g) {
      const Geometry& G = geoms[g];
      data.oMg[g] = G.joint == 0 ? G.placement : compose(data.oMi[G.joint], G.placement);
    }
  }

  // pinocchio computeBodyRadius: per joint, max over its geometries of the largest
  // norm of the 8 corners of the geometry's local AABB mapped by its placement.
  std::vector<double> bodyRadii() const {
    std::vector<double> r(joints.size(), 0.0);
    for (const Geometry& G : geoms) {
      double lo[3], hi[3];
      const Shape& s = G.shape;
      switch (s.type) {
        case SPHERE: for (int k = 0; k < 3; ++k) { lo[k] = -s.p[0]; hi[k] = s.p[0]; } break;
        case CAPSULE: lo[0] = lo[1] = -s.p[0]; hi[0] = hi[1] = s.p[0]; lo[2] = -(s.p[1] + s.p[0]); hi[2] = s.p[1] + s.p[0]; break;
        case BOX: for (int k = 0; k < 3; ++k) { lo[k] = -s.p[k]; hi[k] = s.p[k]; } break;
        default:
          for (int k = 0; k < 3; ++k) { lo[k] = kInf; hi[k] = -kInf; }
          for (int i = 0; i < s.nverts; ++i)
            for (int k = 0; k < 3; ++k) {
              double c = s.verts[3 * i + k];
              if (c < lo[k]) lo[k] = c;
              if (c > hi[k]) hi[k] = c;
            }
      }
      for (int c = 0; c < 8; ++c) {
        Vec3 p{(c & 1) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1], (c & 4) ? hi[2] : lo[2]};
        Vec3 w = mul(G.placement.R, p) + G.placement.t;
        double n = std::sqrt(dot(w, w));
        if (n > r[G.joint]) r[G.joint] = n;
      }
    }
    return r;
  }

  bool jointHasBody(int j) const {
    for (const Geometry& G : geoms) if (G.joint == j) return true;
    return false;
  }

  // hpp-pinocchio Joint::{upperBoundLinearVelocity,upperBoundAngularVelocity,maximalDistanceToParent}
  double upperBoundLinearVelocity(int j) const {
    int t = joints[j].type;
    return (t >= HPPB200_JOINT_PRISMATIC_X && t <= HPPB200_JOINT_FREEFLYER) ? 1.0 : 0.0;
  }
  double upperBoundAngularVelocity(int j) const {
    int t = joints[j].type;
    return (t <= HPPB200_JOINT_REVOLUTE_Z || t == HPPB200_JOINT_FREEFLYER) ? 1.0 : 0.0;
  }
  double maximalDistanceToParent(int j) const {
    const Joint& J = joints[j];
    const Vec3& T = J.placement.t;
    if (J.type <= HPPB200_JOINT_REVOLUTE_Z) return std::sqrt(dot(T, T));
    if (J.type <= HPPB200_JOINT_PRISMATIC_Z) {
      int a = J.type - HPPB200_JOINT_PRISMATIC_X;
      Vec3 u = J.placement.R.col(a);
      Vec3 p0 = axpy(lower[J.idx_q], u, T), p1 = axpy(upper[J.idx_q], u, T);
      double n0 = std::sqrt(dot(p0, p0)), n1 = std::sqrt(dot(p1, p1));
      return n0 > n1 ? n0 : n1;
    }
    // free-flyer: farthest corner of the translation bounds (inf if unbounded)
    double best = 0;
    for (int c = 0; c < 8; ++c) {
      Vec3 p{(c & 1) ? upper[J.idx_q] : lower[J.idx_q], (c & 2) ? upper[J.idx_q + 1] : lower[J.idx_q + 1],
             (c & 4) ? upper[J.idx_q + 2] : lower[J.idx_q + 2]};
      Vec3 w = mul(J.placement.R, p) + T;
      double n = std::sqrt(dot(w, w));
      if (!(n <= best)) best = n;
    }
    return best;
  }
};

}  // namespace oracle
