// hpp-core side of the drop-in boundary: an hpp-core plugin (reference include/hpp/core/plugin.hh:75-80,
// template tests/plugin-test/example.cc:33-47) whose validators and nearest-neighbour object forward to the
// C ABI of include/hppb200.h.  Build it where hpp-core is installed; hpp-core's headers need
// Eigen, Boost, pinocchio and coal, none of which exists in the development image, so this file is NOT part of
// `__graft_entry__.build()`.
//
//   ps->configValidations  "B200CollisionValidation"      -> B200ConfigValidation
//   ps->pathValidations    "B200Discretized" / "B200Progressive" / "B200Dichotomy" -> B200PathValidation
//   ps->pathPlanners       "BiRRTPlanner", "DiffusingPlanner", "VisibilityPrmPlanner" re-registered with
//                          wrappers that install B200NearestNeighbor into the (empty) roadmap first
//                          (reference src/roadmap.cc:74-84: Roadmap::nearestNeighbor(NearestNeighbor*)).
//
// Single-item virtuals (`validate(q)`, `validate(path, ...)`) submit a batch of one; the batch entry points
// `validateBatch` are what batched planner front-ends call.
#include <hppb200.h>

#include <coal/shape/geometric_shapes.h>

#include <hpp/core/collision-validation-report.hh>
#include <hpp/core/config-validation.hh>
#include <hpp/core/connected-component.hh>
#include <hpp/core/distance.hh>
#include <hpp/core/nearest-neighbor.hh>
#include <hpp/core/node.hh>
#include <hpp/core/obstacle-user.hh>
#include <hpp/core/path-validation-report.hh>
#include <hpp/core/path-validation.hh>
#include <hpp/core/plugin.hh>
#include <hpp/core/problem-solver.hh>
#include <hpp/core/problem.hh>
#include <hpp/core/roadmap.hh>
#include <hpp/core/straight-path.hh>
#include <hpp/core/weighed-distance.hh>
#include <hpp/pinocchio/body.hh>
#include <hpp/pinocchio/collision-object.hh>
#include <hpp/pinocchio/device.hh>
#include <hpp/pinocchio/joint.hh>
#include <pinocchio/multibody/geometry.hpp>
#include <pinocchio/multibody/model.hpp>

#include <mutex>
#include <stdexcept>
#include <vector>

namespace hpp {
namespace b200 {
using namespace hpp::core;

static void check(int rc) {
  if (rc == HPPB200_OK) return;
  const std::string msg = hppb200_last_error();
  if (rc == HPPB200_ERR_INVALID_ARGUMENT) throw std::invalid_argument(msg);  // src/obstacle-user.cc:167-175
  throw std::runtime_error(msg);                                             // src/obstacle-user.cc:109
}

static void se3To12(const ::pinocchio::SE3& M, double* out) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) out[3 * i + j] = M.rotation()(i, j);
  for (int i = 0; i < 3; ++i) out[9 + i] = M.translation()[i];
}

// coal::CollisionGeometry -> (HPPB200_SHAPE_*, params, vertices); false for BVH / unsupported shapes
static bool shapeOf(const coal::CollisionGeometry& g, int& type, double p[4], std::vector<double>& verts, int& vb,
                    int& vc) {
  p[0] = p[1] = p[2] = p[3] = 0;
  vb = vc = 0;
  switch (g.getNodeType()) {
    case coal::GEOM_SPHERE: type = HPPB200_SHAPE_SPHERE; p[0] = static_cast<const coal::Sphere&>(g).radius; return true;
    case coal::GEOM_CAPSULE: {
      const auto& c = static_cast<const coal::Capsule&>(g);
      type = HPPB200_SHAPE_CAPSULE; p[0] = c.radius; p[1] = c.halfLength; return true;
    }
    case coal::GEOM_BOX: {
      const auto& b = static_cast<const coal::Box&>(g);
      type = HPPB200_SHAPE_BOX; p[0] = b.halfSide[0]; p[1] = b.halfSide[1]; p[2] = b.halfSide[2]; return true;
    }
    case coal::GEOM_CONVEX: {
      const auto& c = static_cast<const coal::ConvexBase&>(g);
      type = HPPB200_SHAPE_CONVEX; vb = (int)verts.size() / 3; vc = (int)c.num_points;
      for (unsigned i = 0; i < c.num_points; ++i)
        for (int k = 0; k < 3; ++k) verts.push_back((*c.points)[i][k]);
      return true;
    }
    default: return false;
  }
}

// Flattened robot + obstacles (what hppb200_model_create consumes) and the engine handles built from it.
// Rebuilt lazily after every ObstacleUserInterface edit, like ObstacleUser rebuilds cPairs_
// (reference src/obstacle-user.cc:53-241).
class Scene {
 public:
  explicit Scene(const DevicePtr_t& robot) : robot_(robot) {}
  ~Scene() { release(); }
  void addObstacle(const CollisionObjectConstPtr_t& o) { obstacles_.push_back(o); dirty_ = true; }
  void removeObstacleFromJoint(const JointPtr_t&, const CollisionObjectConstPtr_t&) {
    throw std::runtime_error("B200: removeObstacleFromJoint needs per-joint obstacle lists (not yet in the C ABI)");
  }
  void setSecurityMargins(const matrix_t& m) {
    const size_type n = robot_->model().njoints;  // src/obstacle-user.cc:166-175
    if (m.rows() != n || m.cols() != n) throw std::invalid_argument("Wrong size of security margin matrix.");
    margins_.assign(m.data(), m.data() + m.size());
    if (m.IsRowMajor == 0) {  // Eigen default is column major: transpose into the ABI's row-major layout
      for (size_type i = 0; i < n; ++i)
        for (size_type j = 0; j < n; ++j) margins_[i * n + j] = m(i, j);
    }
    dirty_ = true;
  }
  hppb200_workspace* workspace() {  // one workspace (stream + scratch) per calling thread
    std::lock_guard<std::mutex> lock(mu_);
    if (dirty_) rebuild();
    thread_local std::map<const Scene*, std::pair<unsigned, hppb200_workspace*>> tls;
    auto& slot = tls[this];
    if (!slot.second || slot.first != generation_) {
      if (slot.second) hppb200_workspace_destroy(slot.second);
      check(hppb200_workspace_create(model_, 0, &slot.second));
      slot.first = generation_;
    }
    return slot.second;
  }
  hppb200_model* model() { std::lock_guard<std::mutex> lock(mu_); if (dirty_) rebuild(); return model_; }
  const DevicePtr_t& robot() const { return robot_; }
  CollisionObjectConstPtr_t object(int ref) const;  // geometry index (>= 0) or obstacle -(k+1)

 private:
  void release() { if (model_) hppb200_model_destroy(model_); model_ = nullptr; }
  void rebuild();
  DevicePtr_t robot_;
  std::vector<CollisionObjectConstPtr_t> obstacles_;
  std::vector<double> margins_;
  hppb200_model* model_ = nullptr;
  unsigned generation_ = 0;
  bool dirty_ = true;
  std::mutex mu_;
};
typedef shared_ptr<Scene> ScenePtr_t;
