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

// pinocchio::Model / GeometryModel -> hppb200_model_desc (include/hppb200.h:107-146), in pinocchio's own
// order, which is the order ObstacleUser::addRobotCollisionPairs and ContinuousValidation::Initialize walk
// (reference src/obstacle-user.cc:222-235, src/continuous-validation.cc:63-98).
void Scene::rebuild() {
  const ::pinocchio::Model& m = robot_->model();
  const ::pinocchio::GeometryModel& gm = robot_->geomModel();
  const int nj = m.njoints - 1;
  std::vector<int32_t> parent(nj), type(nj), gjoint, gshape, gvb, gvc, self, oshape, ovb, ovc;
  std::vector<double> jplace(12 * nj), lower(m.nq), upper(m.nq), gplace, gparams, oplace, oparams, verts, weights(nj, 1.0);
  for (int j = 1; j <= nj; ++j) {
    parent[j - 1] = (int)m.parents[j];
    const std::string s = m.joints[j].shortname();
    if (s == "JointModelRX") type[j - 1] = HPPB200_JOINT_REVOLUTE_X;
    else if (s == "JointModelRY") type[j - 1] = HPPB200_JOINT_REVOLUTE_Y;
    else if (s == "JointModelRZ") type[j - 1] = HPPB200_JOINT_REVOLUTE_Z;
    else if (s == "JointModelPX") type[j - 1] = HPPB200_JOINT_PRISMATIC_X;
    else if (s == "JointModelPY") type[j - 1] = HPPB200_JOINT_PRISMATIC_Y;
    else if (s == "JointModelPZ") type[j - 1] = HPPB200_JOINT_PRISMATIC_Z;
    else if (s == "JointModelFreeFlyer") type[j - 1] = HPPB200_JOINT_FREEFLYER;
    else throw std::runtime_error("B200: unsupported joint model " + s);
    se3To12(m.jointPlacements[j], &jplace[12 * (j - 1)]);
  }
  for (int k = 0; k < m.nq; ++k) { lower[k] = m.lowerPositionLimit[k]; upper[k] = m.upperPositionLimit[k]; }
  for (const ::pinocchio::GeometryObject& g : gm.geometryObjects) {
    int t, vb, vc; double p[4];
    if (!shapeOf(*g.geometry, t, p, verts, vb, vc))
      throw std::runtime_error("B200: geometry '" + g.name + "' is a mesh/BVH; convexify it first (DESIGN.md 9)");
    gjoint.push_back((int)g.parentJoint); gshape.push_back(t); gvb.push_back(vb); gvc.push_back(vc);
    gparams.insert(gparams.end(), p, p + 4);
    gplace.resize(gplace.size() + 12); se3To12(g.placement, &gplace[gplace.size() - 12]);
  }
  for (std::size_t k = 0; k < gm.collisionPairs.size(); ++k) {  // active pairs only, in order
    if (!robot_->geomData().activeCollisionPairs[k]) continue;
    self.push_back((int)gm.collisionPairs[k].first); self.push_back((int)gm.collisionPairs[k].second);
  }
  for (const CollisionObjectConstPtr_t& o : obstacles_) {
    int t, vb, vc; double p[4];
    if (!shapeOf(*o->geometry(), t, p, verts, vb, vc))
      throw std::runtime_error("B200: obstacle '" + o->name() + "' is a mesh/BVH; convexify it first");
    oshape.push_back(t); ovb.push_back(vb); ovc.push_back(vc);
    oparams.insert(oparams.end(), p, p + 4);
    oplace.resize(oplace.size() + 12); se3To12(o->getTransform(), &oplace[oplace.size() - 12]);
  }
  hppb200_model_desc d{};
  d.njoints = nj; d.joint_parent = parent.data(); d.joint_type = type.data(); d.joint_placement = jplace.data();
  d.q_lower = lower.data(); d.q_upper = upper.data();
  d.ngeoms = (int)gjoint.size(); d.geom_joint = gjoint.data(); d.geom_placement = gplace.data();
  d.geom_shape = gshape.data(); d.geom_params = gparams.data(); d.geom_vert_begin = gvb.data(); d.geom_vert_count = gvc.data();
  d.nself_pairs = (int)self.size() / 2; d.self_pairs = self.data();
  d.nobstacles = (int)oshape.size(); d.obst_placement = oplace.data(); d.obst_shape = oshape.data();
  d.obst_params = oparams.data(); d.obst_vert_begin = ovb.data(); d.obst_vert_count = ovc.data();
  d.nvertices = (int)verts.size() / 3; d.vertices = verts.data();
  d.security_margins = margins_.empty() ? nullptr : margins_.data();
  d.break_distance = 1e-2;
  d.weights = weights.data();  // replaced by WeighedDistance::weights() when a distance is attached
  release();
  check(hppb200_model_create(&d, &model_));
  ++generation_;
  dirty_ = false;
}

CollisionObjectConstPtr_t Scene::object(int ref) const {
  if (ref < 0) return obstacles_[-(ref + 1)];
  return CollisionObjectConstPtr_t(new pinocchio::CollisionObject(robot_, (std::size_t)ref));
}

// CollisionValidationReport for pair `index` of table (geomA, geomB) (reference collision-validation-report.hh:55-66)
static ValidationReportPtr_t collisionReport(Scene& s, const std::vector<int32_t>& A, const std::vector<int32_t>& B, int index) {
  CollisionValidationReportPtr_t r(new CollisionValidationReport);
  r->object1 = s.object(A[index]);
  r->object2 = s.object(B[index]);
  return r;
}

// ---- ConfigValidation -------------------------------------------------------------------------------
class B200ConfigValidation : public ConfigValidation, public ObstacleUserInterface {
 public:
  static shared_ptr<B200ConfigValidation> create(const DevicePtr_t& robot) {
    return shared_ptr<B200ConfigValidation>(new B200ConfigValidation(robot));
  }
  // reference include/hpp/core/config-validation.hh:55-56
  bool validate(ConfigurationIn_t q, ValidationReportPtr_t& report) override {
    uint8_t valid, reason; int32_t first;
    check(hppb200_validate_configs(scene_->workspace(), q.data(), 1, 0, &valid, &first, &reason));
    if (valid) return true;
    std::vector<int32_t> A(hppb200_model_num_config_pairs(scene_->model())), B(A.size());
    hppb200_model_get_config_pairs(scene_->model(), A.data(), B.data(), nullptr);
    report = collisionReport(*scene_, A, B, first);
    return false;
  }
  // batch extension: q row-major [n][nq]
  void validateBatch(const double* q, std::size_t n, uint8_t* valid, int32_t* firstPair) {
    check(hppb200_validate_configs(scene_->workspace(), q, n, 0, valid, firstPair, nullptr));
  }
  void addObstacle(const CollisionObjectConstPtr_t& o) override { scene_->addObstacle(o); }
  void removeObstacleFromJoint(const JointPtr_t& j, const CollisionObjectConstPtr_t& o) override { scene_->removeObstacleFromJoint(j, o); }
  void filterCollisionPairs(const RelativeMotion::matrix_type&) override {}  // constraints-driven; none on this path
  void setSecurityMargins(const matrix_t& m) override { scene_->setSecurityMargins(m); }
  void setSecurityMarginBetweenBodies(const std::string&, const std::string&, const value_type&) override {
    throw std::invalid_argument("B200: per-body margins go through setSecurityMargins");
  }
  const ScenePtr_t& scene() const { return scene_; }

 private:
  explicit B200ConfigValidation(const DevicePtr_t& robot) : scene_(new Scene(robot)) {}
  ScenePtr_t scene_;
};

// ---- PathValidation ---------------------------------------------------------------------------------
// Replaces pathValidation::Discretized (src/path-validation/discretized.cc:55-107),
// continuousValidation::Progressive (progressive.cc:78-129) and ::Dichotomy (dichotomy.cc:68-132) for
// StraightPath edges; any other path type is forwarded to the stock validator given at construction.
class B200PathValidation : public PathValidation, public ObstacleUserInterface {
 public:
  static PathValidationPtr_t create(const DevicePtr_t& robot, int method, value_type param, PathValidationPtr_t stock) {
    // constructor checks of the reference: progressive.cc:138-145, dichotomy.cc:139-148
    if (method == HPPB200_METHOD_PROGRESSIVE && param <= 0)
      throw std::runtime_error("tolerance should be positive for progressive continuous validation.");
    if (method == HPPB200_METHOD_DICHOTOMY && param != 0)
      throw std::runtime_error("Dichotomy path validation method does not support penetration.");
    return PathValidationPtr_t(new B200PathValidation(robot, method, param, stock));
  }
  // reference include/hpp/core/path-validation.hh:62-64
  bool validate(const PathPtr_t& path, bool reverse, PathPtr_t& validPart, PathValidationReportPtr_t& report) override {
    StraightPathPtr_t sp = HPP_DYNAMIC_PTR_CAST(StraightPath, path);
    if (!sp || sp->constraints()) return stock_->validate(path, reverse, validPart, report);
    const interval_t tr = path->timeRange();
    const Configuration_t q1 = sp->initial(), q2 = sp->end();
    const double L = tr.second - tr.first;
    uint8_t valid, reason; int32_t pair; double tFail, tLast;
    check(hppb200_validate_edges(scene_->workspace(), method_, param_, reverse ? 1 : 0, q1.data(), q2.data(), &L, 1, 0,
                                 &valid, &tFail, &tLast, &pair, &reason));
    if (valid) { validPart = path; return true; }
    // ownership rule: validPart is set on every return (progressive.cc:121-127, discretized.cc:99-105)
    validPart = reverse ? path->extract(tr.first + tLast, tr.second) : path->extract(tr.first, tr.first + tLast);
    hppb200_model* m = scene_->model();
    const bool disc = method_ == HPPB200_METHOD_DISCRETIZED;
    std::vector<int32_t> A(disc ? hppb200_model_num_config_pairs(m) : hppb200_model_num_body_geom_pairs(m)), B(A.size());
    if (disc) hppb200_model_get_config_pairs(m, A.data(), B.data(), nullptr);
    else hppb200_model_get_body_geom_pairs(m, A.data(), B.data(), nullptr);
    report = PathValidationReportPtr_t(new PathValidationReport(tr.first + tFail, collisionReport(*scene_, A, B, pair)));
    return false;
  }
  // batch extension: edges as (q1, q2, L) arrays; see include/hppb200.h:233-243
  void validateBatch(const double* q1, const double* q2, const double* L, std::size_t n, bool reverse, uint8_t* valid,
                     double* tFail, double* tLast, int32_t* pair) {
    check(hppb200_validate_edges(scene_->workspace(), method_, param_, reverse ? 1 : 0, q1, q2, L, n, 0, valid, tFail,
                                 tLast, pair, nullptr));
  }
  void addObstacle(const CollisionObjectConstPtr_t& o) override { scene_->addObstacle(o); forward([&](ObstacleUserInterface& u) { u.addObstacle(o); }); }
  void removeObstacleFromJoint(const JointPtr_t& j, const CollisionObjectConstPtr_t& o) override { scene_->removeObstacleFromJoint(j, o); }
  void filterCollisionPairs(const RelativeMotion::matrix_type& r) override { forward([&](ObstacleUserInterface& u) { u.filterCollisionPairs(r); }); }
  void setSecurityMargins(const matrix_t& m) override { scene_->setSecurityMargins(m); forward([&](ObstacleUserInterface& u) { u.setSecurityMargins(m); }); }
  void setSecurityMarginBetweenBodies(const std::string& a, const std::string& b, const value_type& v) override {
    forward([&](ObstacleUserInterface& u) { u.setSecurityMarginBetweenBodies(a, b, v); });
    throw std::invalid_argument("B200: per-body margins go through setSecurityMargins");
  }

 private:
  B200PathValidation(const DevicePtr_t& robot, int method, value_type param, PathValidationPtr_t stock)
      : scene_(new Scene(robot)), method_(method), param_(param), stock_(stock) {}
  template <class F> void forward(F f) {
    shared_ptr<ObstacleUserInterface> u = HPP_DYNAMIC_PTR_CAST(ObstacleUserInterface, stock_);
    if (u) f(*u);
  }
  ScenePtr_t scene_;
  int method_;
  value_type param_;
  PathValidationPtr_t stock_;
};

// ---- NearestNeighbor --------------------------------------------------------------------------------
// Replaces nearestNeighbor::Basic (src/nearest-neighbor/basic.cc:52-181): nodes are mirrored on the device in
// insertion order (the reference's scan order), connected components as integer labels.
class B200NearestNeighbor : public NearestNeighbor {
 public:
  B200NearestNeighbor(const DistancePtr_t& d, const ScenePtr_t& scene) : distance_(d), scene_(scene) {
    check(hppb200_knn_create(scene_->model(), 0, &nn_));
  }
  ~B200NearestNeighbor() override { hppb200_knn_destroy(nn_); }
  void clear() override { hppb200_knn_destroy(nn_); nn_ = nullptr; check(hppb200_knn_create(scene_->model(), 0, &nn_)); nodes_.clear(); labels_.clear(); }
  void addNode(const NodePtr_t& node) override {
    const int32_t cc = label(node->connectedComponent());
    check(hppb200_knn_append(nn_, node->configuration().data(), &cc, 1));
    nodes_.push_back(node);
  }
  NodePtr_t search(ConfigurationIn_t q, const ConnectedComponentPtr_t& cc, value_type& d, bool) override {
    int64_t idx; check(hppb200_knn_query(nn_, q.data(), 1, 1, label(cc), 0, &idx, &d));
    return idx < 0 ? NodePtr_t() : nodes_[(std::size_t)idx];  // WeighedDistance is symmetric: `reverse` is a no-op
  }
  NodePtr_t search(const NodePtr_t& n, const ConnectedComponentPtr_t& cc, value_type& d) override {
    return search(n->configuration(), cc, d, false);
  }
  Nodes_t KnearestSearch(ConfigurationIn_t q, const ConnectedComponentPtr_t& cc, const std::size_t K, value_type& d) override {
    return knn(q, label(cc), K, d);
  }
  Nodes_t KnearestSearch(const NodePtr_t& n, const ConnectedComponentPtr_t& cc, const std::size_t K, value_type& d) override {
    return knn(n->configuration(), label(cc), K, d);
  }
  Nodes_t KnearestSearch(ConfigurationIn_t q, const RoadmapPtr_t&, const std::size_t K, value_type& d) override {
    return knn(q, -1, K, d);
  }
  NodeVector_t withinBall(ConfigurationIn_t q, const ConnectedComponentPtr_t& cc, value_type maxD) override {
    std::vector<int64_t> idx(nodes_.size()); int64_t count = 0;
    check(hppb200_knn_within_ball(nn_, q.data(), 1, maxD, label(cc), idx.size(), idx.data(), &count));
    NodeVector_t out;
    for (int64_t i = 0; i < count; ++i) out.push_back(nodes_[(std::size_t)idx[i]]);
    return out;
  }
  void merge(ConnectedComponentPtr_t cc1, ConnectedComponentPtr_t cc2) override {
    check(hppb200_knn_relabel(nn_, label(cc2), label(cc1)));
    labels_[cc2.get()] = label(cc1);
  }
  DistancePtr_t distance() const override { return distance_; }

 private:
  int32_t label(const ConnectedComponentPtr_t& cc) {
    auto it = labels_.find(cc.get());
    if (it == labels_.end()) it = labels_.insert({cc.get(), (int32_t)labels_.size()}).first;
    return it->second;
  }
  Nodes_t knn(ConfigurationIn_t q, int32_t cc, std::size_t K, value_type& d) {
    std::vector<int64_t> idx(K); std::vector<double> dist(K);
    check(hppb200_knn_query(nn_, q.data(), 1, (int)K, cc, 0, idx.data(), dist.data()));
    Nodes_t out;
    for (std::size_t i = 0; i < K && idx[i] >= 0; ++i) { out.push_back(nodes_[(std::size_t)idx[i]]); d = dist[i]; }
    return out;
  }
  DistancePtr_t distance_;
  ScenePtr_t scene_;
  hppb200_knn* nn_ = nullptr;
  std::vector<NodePtr_t> nodes_;
  std::map<const ConnectedComponent*, int32_t> labels_;
};

// ---- registration (reference src/plugin.cc:43-104, include/hpp/core/problem-solver.hh:52-59,519-522) ----
class B200Plugin : public ProblemSolverPlugin {
 public:
  B200Plugin() : ProblemSolverPlugin("B200ValidationPlugin", "0.1") {}

 protected:
  bool impl_initialize(ProblemSolverPtr_t ps) override {
    ps->configValidations.add("B200CollisionValidation", [](const DevicePtr_t& r) -> ConfigValidationPtr_t { return B200ConfigValidation::create(r); });
    struct M { const char* name; const char* stock; int method; };
    for (const M& m : {M{"B200Discretized", "Discretized", HPPB200_METHOD_DISCRETIZED},
                       M{"B200Progressive", "Progressive", HPPB200_METHOD_PROGRESSIVE},
                       M{"B200Dichotomy", "Dichotomy", HPPB200_METHOD_DICHOTOMY}}) {
      PathValidationBuilder_t stock = ps->pathValidations.get(m.stock);
      const int method = m.method;
      ps->pathValidations.add(m.name, [stock, method](const DevicePtr_t& r, const value_type& p) {
        return B200PathValidation::create(r, method, p, stock(r, p));
      });
    }
    for (const char* planner : {"BiRRTPlanner", "DiffusingPlanner", "VisibilityPrmPlanner"}) {
      PathPlannerBuilder_t stock = ps->pathPlanners.get(planner);
      ps->pathPlanners.add(planner, [stock](const ProblemConstPtr_t& problem, const RoadmapPtr_t& roadmap) {
        if (roadmap->nodes().empty()) {  // Roadmap::nearestNeighbor requires an empty roadmap (src/roadmap.cc:74-84)
          ScenePtr_t scene(new Scene(problem->robot()));
          roadmap->nearestNeighbor(new B200NearestNeighbor(problem->distance(), scene));
        }
        return stock(problem, roadmap);
      });
    }
    return true;
  }
};
}  // namespace b200
}  // namespace hpp

HPP_CORE_DEFINE_PLUGIN(hpp::b200::B200Plugin)
