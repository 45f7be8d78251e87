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
// `validateBatch` are what batched planner front-ends call.  A PathVector of StraightPaths goes to
// hppb200_validate_edge_chains (one batch, cut at the first failing leaf); other path types to the stock validator.
// Everything that does not need hpp-core's headers (component labels, workspace pool) lives in b200_adapter_util.hh
// and is tested with mocks (tests/plugin_mock/).
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
#include <hpp/core/joint-bound-validation.hh>
#include <hpp/core/path-validation.hh>
#include <hpp/core/path-vector.hh>
#include <hpp/core/path/spline.hh>
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
#include <hpp/util/serialization.hh>
#include <pinocchio/multibody/geometry.hpp>
#include <pinocchio/multibody/model.hpp>

#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "b200_adapter_util.hh"

namespace hpp {
namespace b200 {
using namespace hpp::core;

static void check(int rc) {
  if (rc == HPPB200_OK) return;
  const std::string msg = hppb200_last_error();
  if (rc == HPPB200_ERR_INVALID_ARGUMENT) throw std::invalid_argument(msg);  // src/obstacle-user.cc:167-175, :216-219
  throw std::runtime_error(msg);                                             // src/obstacle-user.cc:104-110
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

// Flattened robot + obstacles (what hppb200_model_create consumes), the LOG of ObstacleUserInterface calls received
// after construction (hppb200_pair_op, include/hppb200.h) and the engine handles built from them.  Rebuilt lazily
// after every edit, like ObstacleUser edits cPairs_ in place (reference src/obstacle-user.cc:53-241).
// Every validator owns its Scene (its own pair tables and call log, like every ObstacleUser in the reference); the
// nearest-neighbour object borrows the Scene of the problem's path validation when that is one of ours (one model
// upload instead of two) and only reads the robot part and the weights.  As in the reference (the pool of
// BodyPairCollision copies is cleared by every edit, src/continuous-validation.cc:243,299), edits must not overlap
// validate calls.
class Scene {
 public:
  typedef WorkspacePool<hppb200_workspace> Pool;
  explicit Scene(const DevicePtr_t& robot)
      : robot_(robot),
        pool_([this]() { hppb200_workspace* w = nullptr; check(hppb200_workspace_create(model_, 0, &w)); return w; },
              [](hppb200_workspace* w) { hppb200_workspace_destroy(w); }) {}
  ~Scene() { release(); }

  // ---- ObstacleUserInterface, recorded (reference include/hpp/core/obstacle-user.hh:47-88) ----
  void addObstacle(const CollisionObjectConstPtr_t& o) {
    std::lock_guard<std::mutex> lock(mu_);
    obstacles_.push_back(o);
    // before any other edit: part of the base tables; after: an op, so that earlier filters / margins do not apply
    const bool late = !ops_.empty();
    obstGlobal_.push_back(late ? 0 : 1);
    if (late) push(HPPB200_OP_ADD_OBSTACLE, (int)obstacles_.size() - 1, 0, 0, 0.0);
    dirty_ = true;
  }
  void removeObstacleFromJoint(const JointPtr_t& joint, const CollisionObjectConstPtr_t& o) {
    std::lock_guard<std::mutex> lock(mu_);
    const int k = obstacleIndex(o);
    if (k < 0)  // src/obstacle-user.cc:104-110
      throw std::runtime_error("ObstacleUser::removeObstacleFromJoint: obstacle \"" + o->name() +
                               "\" is not registered as obstacle for joint \"" + joint->name() + "\".");
    push(HPPB200_OP_REMOVE_OBSTACLE_FROM_JOINT, k, (int)joint->index(), 0, 0.0);
    commitOrRollback();
  }
  void filterCollisionPairs(const RelativeMotion::matrix_type& m) {
    std::lock_guard<std::mutex> lock(mu_);
    std::vector<int32_t> rel((std::size_t)(m.rows() * m.cols()));
    for (size_type i = 0; i < m.rows(); ++i)
      for (size_type j = 0; j < m.cols(); ++j) rel[(std::size_t)(i * m.cols() + j)] = (int32_t)m(i, j);  // 0 Constrained, 1 Parameterized, 2 Unconstrained
    intData_.push_back(std::move(rel));
    push(HPPB200_OP_FILTER_COLLISION_PAIRS, 0, 0, 0, 0.0, intData_.back().data());
    commitOrRollback();
  }
  void setSecurityMargins(const matrix_t& m) {
    std::lock_guard<std::mutex> lock(mu_);
    const size_type n = robot_->nbJoints() + 1;  // src/obstacle-user.cc:166-175
    if (m.rows() != n || m.cols() != n) throw std::invalid_argument("Wrong size of security margin matrix.");
    std::vector<double> mat((std::size_t)(n * n));
    for (size_type i = 0; i < n; ++i)
      for (size_type j = 0; j < n; ++j) mat[(std::size_t)(i * n + j)] = m(i, j);  // row major, whatever Eigen's storage order
    dblData_.push_back(std::move(mat));
    push(HPPB200_OP_SET_SECURITY_MARGINS, 0, 0, 0, 0.0, dblData_.back().data());
    commitOrRollback();
  }
  void setSecurityMarginBetweenBodies(const std::string& a, const std::string& b, const value_type& margin) {
    std::lock_guard<std::mutex> lock(mu_);
    int ra, rb;
    if (!bodyRef(a, ra) || !bodyRef(b, rb))  // src/obstacle-user.cc:216-219
      throw std::invalid_argument("Could not find a collision pair between body " + a + " and " + b);
    push(HPPB200_OP_SET_MARGIN_BETWEEN_BODIES, ra, rb, 0, margin);
    commitOrRollback();
  }
  void checkParameterized(bool on) { std::lock_guard<std::mutex> lock(mu_); checkParameterized_ = on; dirty_ = true; }
  // WeighedDistance::weights() of the problem's distance (include/hpp/core/weighed-distance.hh:60); false if the
  // distance is not a WeighedDistance over this robot: the caller then keeps the stock nearest-neighbour object
  bool setDistance(const DistancePtr_t& d) {
    WeighedDistancePtr_t wd = HPP_DYNAMIC_PTR_CAST(WeighedDistance, d);
    if (!wd || wd->size() != robot_->nbJoints()) return false;
    std::lock_guard<std::mutex> lock(mu_);
    std::vector<double> w((std::size_t)wd->size());
    for (size_type i = 0; i < wd->size(); ++i) w[(std::size_t)i] = wd->weights()[i];
    if (w != weights_) { weights_ = w; dirty_ = true; }
    return true;
  }

  // ---- engine handles ----
  // Borrow a workspace (stream + scratch) for one call; concurrent validate calls get distinct workspaces
  // (tests/test-continuous-validation.cc:324-343 calls validate from an OpenMP loop).
  Pool::Lease workspace() {
    std::lock_guard<std::mutex> lock(mu_);
    if (dirty_) rebuild();
    return pool_.acquire();
  }
  hppb200_model* model() { std::lock_guard<std::mutex> lock(mu_); if (dirty_) rebuild(); return model_; }
  const DevicePtr_t& robot() const { return robot_; }
  CollisionObjectConstPtr_t object(int ref) const;  // geometry index (>= 0) or obstacle -(k+1)

 private:
  void push(int kind, int a, int b, int flag, double value, const void* data = nullptr) {
    hppb200_pair_op op{};
    op.kind = kind; op.a = a; op.b = b; op.flag = flag; op.value = value; op.data = data;
    ops_.push_back(op);
    dirty_ = true;
  }
  // the reference throws from the edit itself (unknown obstacle / body pair): build now, drop the op if it is rejected
  void commitOrRollback() {
    try { rebuild(); } catch (...) { ops_.pop_back(); dirty_ = true; throw; }
  }
  int obstacleIndex(const CollisionObjectConstPtr_t& o) const {
    for (std::size_t k = 0; k < obstacles_.size(); ++k)
      if (obstacles_[k] == o || obstacles_[k]->name() == o->name()) return (int)k;
    return -1;
  }
  bool bodyRef(const std::string& name, int& ref) const {
    const ::pinocchio::GeometryModel& gm = robot_->geomModel();
    for (std::size_t g = 0; g < gm.geometryObjects.size(); ++g)
      if (gm.geometryObjects[g].name == name) { ref = (int)g; return true; }
    for (std::size_t k = 0; k < obstacles_.size(); ++k)
      if (obstacles_[k]->name() == name) { ref = -((int)k + 1); return true; }
    return false;
  }
  void release() {
    pool_.clear();  // workspaces first: they point into the model
    if (model_) hppb200_model_destroy(model_);
    model_ = nullptr;
  }
  void rebuild();
  DevicePtr_t robot_;
  std::vector<CollisionObjectConstPtr_t> obstacles_;
  std::vector<uint8_t> obstGlobal_;
  std::vector<hppb200_pair_op> ops_;
  std::vector<std::vector<int32_t>> intData_;  // matrices the ops point to (kept for the life of the scene)
  std::vector<std::vector<double>> dblData_;
  std::vector<double> weights_;
  bool checkParameterized_ = false;
  hppb200_model* model_ = nullptr;
  bool dirty_ = true;
  std::mutex mu_;
  Pool pool_;  // declared last: destroyed first
};
typedef shared_ptr<Scene> ScenePtr_t;


// pinocchio::Model / GeometryModel -> hppb200_model_desc (include/hppb200.h), in pinocchio's own order, which is the
// order ObstacleUser::addRobotCollisionPairs and ContinuousValidation::Initialize walk
// (reference src/obstacle-user.cc:222-235, src/continuous-validation.cc:63-98).  Called with mu_ held.
void Scene::rebuild() {
  const ::pinocchio::Model& m = robot_->model();
  const ::pinocchio::GeometryModel& gm = robot_->geomModel();
  const int nj = m.njoints - 1;
  std::vector<int32_t> parent(nj), type(nj), gjoint, gshape, gvb, gvc, self, oshape, ovb, ovc;
  std::vector<double> jplace(12 * nj), lower(m.nq), upper(m.nq), gplace, gparams, oplace, oparams, verts;
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
  std::vector<double> weights(weights_);
  if (weights.empty()) weights.assign(nj, 1.0);  // no distance attached yet: only the nearest-neighbour object reads them
  hppb200_model_desc d{};
  d.njoints = nj; d.joint_parent = parent.data(); d.joint_type = type.data(); d.joint_placement = jplace.data();
  d.q_lower = lower.data(); d.q_upper = upper.data();
  d.ngeoms = (int)gjoint.size(); d.geom_joint = gjoint.data(); d.geom_placement = gplace.data();
  d.geom_shape = gshape.data(); d.geom_params = gparams.data(); d.geom_vert_begin = gvb.data(); d.geom_vert_count = gvc.data();
  d.nself_pairs = (int)self.size() / 2; d.self_pairs = self.data();
  d.nobstacles = (int)oshape.size(); d.obst_placement = oplace.data(); d.obst_shape = oshape.data();
  d.obst_params = oparams.data(); d.obst_vert_begin = ovb.data(); d.obst_vert_count = ovc.data();
  d.nvertices = (int)verts.size() / 3; d.vertices = verts.data();
  d.security_margins = nullptr;  // margins always arrive as ops: a matrix only applies to the pairs that exist when it is set
  d.break_distance = 1e-2;       // ContinuousValidation::breakDistance_ default (src/continuous-validation.cc:411)
  d.weights = weights.data();
  d.obst_global = obstGlobal_.empty() ? nullptr : obstGlobal_.data();
  d.npair_ops = (int)ops_.size();
  d.pair_ops = ops_.empty() ? nullptr : ops_.data();
  d.check_parameterized = checkParameterized_ ? 1 : 0;
  hppb200_model* fresh = nullptr;
  check(hppb200_model_create(&d, &fresh));  // throws (std::invalid_argument / std::runtime_error) and keeps the old model
  release();
  model_ = fresh;
  dirty_ = false;
}

CollisionObjectConstPtr_t Scene::object(int ref) const {
  if (ref < 0) return obstacles_[-(ref + 1)];
  return CollisionObjectConstPtr_t(new pinocchio::CollisionObject(robot_, (std::size_t)ref));
}

// CollisionValidationReport for pair `index` of table (geomA, geomB) (reference collision-validation-report.hh:42-66).
// `result` stays empty: the engine's narrow phase produces the decision and the distance, not coal's contact patch;
// callers that read result.getContact() must run the stock validator on the reported configuration.
static ValidationReportPtr_t collisionReport(Scene& s, const std::vector<int32_t>& A, const std::vector<int32_t>& B, int index) {
  CollisionValidationReportPtr_t r(new CollisionValidationReport);
  if (index < 0 || index >= (int)A.size()) return r;  // (e.g. HPPB200_REASON_ITER_CAP carries no pair)
  r->object1 = s.object(A[index]);
  r->object2 = s.object(B[index]);
  r->objectName1 = r->object1->name();
  r->objectName2 = r->object2->name();
  return r;
}

// JointBoundValidationReport for configuration rank `rank` (reference src/joint-bound-validation.cc:49-62)
static ValidationReportPtr_t jointBoundReport(const DevicePtr_t& robot, int rank, value_type value) {
  const ::pinocchio::Model& m = robot->model();
  JointPtr_t joint = robot->getJointAtConfigRank((size_type)rank);
  return ValidationReportPtr_t(new JointBoundValidationReport(joint, (size_type)rank - joint->rankInConfiguration(),
                                                              m.lowerPositionLimit[rank], m.upperPositionLimit[rank], value));
}

// ---- ConfigValidation -------------------------------------------------------------------------------
// "B200CollisionValidation": CollisionValidation (src/collision-validation.cc:48-95).
// "B200ConfigValidations": JointBoundValidation first, then collision, in one device pass
// (src/config-validations.cc:40-49 with the default Problem's two validators, src/joint-bound-validation.cc:45-80).
class B200ConfigValidation : public ConfigValidation, public ObstacleUserInterface {
 public:
  static shared_ptr<B200ConfigValidation> create(const DevicePtr_t& robot, bool checkBounds) {
    return shared_ptr<B200ConfigValidation>(new B200ConfigValidation(robot, checkBounds));
  }
  // reference include/hpp/core/config-validation.hh:55-56
  bool validate(ConfigurationIn_t q, ValidationReportPtr_t& report) override {
    uint8_t valid, reason; int32_t first;
    {
      Scene::Pool::Lease ws = scene_->workspace();
      check(hppb200_validate_configs(ws.get(), q.data(), 1, flags_, &valid, &first, &reason));
    }
    if (valid) return true;
    if (reason == HPPB200_REASON_JOINT_BOUND) { report = jointBoundReport(scene_->robot(), first, q[first]); return false; }
    std::vector<int32_t> A(hppb200_model_num_config_pairs(scene_->model())), B(A.size());
    hppb200_model_get_config_pairs(scene_->model(), A.data(), B.data(), nullptr);
    report = collisionReport(*scene_, A, B, first);
    return false;
  }
  // batch extension: q row-major [n][nq]; reason HPPB200_REASON_* (may be null)
  void validateBatch(const double* q, std::size_t n, uint8_t* valid, int32_t* firstPair, uint8_t* reason = nullptr) {
    Scene::Pool::Lease ws = scene_->workspace();
    check(hppb200_validate_configs(ws.get(), q, n, flags_, valid, firstPair, reason));
  }
  /// CollisionValidation::checkParameterized (include/hpp/core/collision-validation.hh)
  void checkParameterized(bool on) { scene_->checkParameterized(on); }
  void addObstacle(const CollisionObjectConstPtr_t& o) override { scene_->addObstacle(o); }
  void removeObstacleFromJoint(const JointPtr_t& j, const CollisionObjectConstPtr_t& o) override { scene_->removeObstacleFromJoint(j, o); }
  void filterCollisionPairs(const RelativeMotion::matrix_type& m) override { scene_->filterCollisionPairs(m); }
  void setSecurityMargins(const matrix_t& m) override { scene_->setSecurityMargins(m); }
  void setSecurityMarginBetweenBodies(const std::string& a, const std::string& b, const value_type& v) override {
    scene_->setSecurityMarginBetweenBodies(a, b, v);
  }
  const ScenePtr_t& scene() const { return scene_; }

 private:
  B200ConfigValidation(const DevicePtr_t& robot, bool checkBounds)
      : scene_(new Scene(robot)), flags_(checkBounds ? HPPB200_CHECK_JOINT_BOUNDS : 0u) {}
  ScenePtr_t scene_;
  uint32_t flags_;
};

// ---- PathValidation ---------------------------------------------------------------------------------
// Replaces pathValidation::Discretized (src/path-validation/discretized.cc:55-107),
// continuousValidation::Progressive (progressive.cc:78-129) and ::Dichotomy (dichotomy.cc:68-132) for
// StraightPath edges, and the PathVector recursion of ContinuousValidation::validate
// (src/continuous-validation.cc:153-192) for PathVectors whose leaves are all unconstrained StraightPaths; any other
// path is forwarded to the stock validator given at construction, which receives the same ObstacleUserInterface calls.
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
    if (sp && !sp->constraints()) return validateStraight(sp, reverse, validPart, report);
    PathVectorPtr_t pv = HPP_DYNAMIC_PTR_CAST(PathVector, path);
    if (pv && method_ != HPPB200_METHOD_DISCRETIZED) {
      std::vector<StraightPathPtr_t> leaves;
      if (flatten(pv, leaves)) return validateChain(pv, leaves, reverse, validPart, report);
    }
    // cubic Bezier splines (steeringMethod::Spline<BernsteinBasis, 3>, src/steering-method/spline.cc) on robots whose
    // configuration space is a vector space: hppb200_validate_spline_edges
    typedef path::Spline<path::BernsteinBasis, 3> Spline3_t;
    shared_ptr<Spline3_t> spl = HPP_DYNAMIC_PTR_CAST(Spline3_t, path);
    if (spl && !spl->constraints() && scene_->robot()->configSize() == scene_->robot()->numberDof())
      return validateSpline(spl, reverse, validPart, report);
    return stock_->validate(path, reverse, validPart, report);
  }
  // batch extension: edges as (q1, q2, L) arrays; see hppb200_validate_edges in include/hppb200.h
  void validateBatch(const double* q1, const double* q2, const double* L, std::size_t n, bool reverse, uint8_t* valid,
                     double* tFail, double* tLast, int32_t* pair, uint8_t* reason = nullptr) {
    Scene::Pool::Lease ws = scene_->workspace();
    check(hppb200_validate_edges(ws.get(), method_, param_, reverse ? 1 : 0, q1, q2, L, n, 0, valid, tFail, tLast, pair, reason));
  }
  void addObstacle(const CollisionObjectConstPtr_t& o) override { scene_->addObstacle(o); forward([&](ObstacleUserInterface& u) { u.addObstacle(o); }); }
  void removeObstacleFromJoint(const JointPtr_t& j, const CollisionObjectConstPtr_t& o) override {
    scene_->removeObstacleFromJoint(j, o);
    forward([&](ObstacleUserInterface& u) { u.removeObstacleFromJoint(j, o); });
  }
  void filterCollisionPairs(const RelativeMotion::matrix_type& r) override {
    scene_->filterCollisionPairs(r);
    forward([&](ObstacleUserInterface& u) { u.filterCollisionPairs(r); });
  }
  void setSecurityMargins(const matrix_t& m) override { scene_->setSecurityMargins(m); forward([&](ObstacleUserInterface& u) { u.setSecurityMargins(m); }); }
  void setSecurityMarginBetweenBodies(const std::string& a, const std::string& b, const value_type& v) override {
    scene_->setSecurityMarginBetweenBodies(a, b, v);
    forward([&](ObstacleUserInterface& u) { u.setSecurityMarginBetweenBodies(a, b, v); });
  }
  const ScenePtr_t& scene() const { return scene_; }

 private:
  B200PathValidation(const DevicePtr_t& robot, int method, value_type param, PathValidationPtr_t stock)
      : scene_(new Scene(robot)), method_(method), param_(param), stock_(stock) {}
  template <class F> void forward(F f) {
    shared_ptr<ObstacleUserInterface> u = HPP_DYNAMIC_PTR_CAST(ObstacleUserInterface, stock_);
    if (u) f(*u);
  }
  // all leaves of the (possibly nested) PathVector, in order; false if one of them is not an unconstrained StraightPath
  static bool flatten(const PathVectorPtr_t& pv, std::vector<StraightPathPtr_t>& out) {
    for (std::size_t i = 0; i < pv->numberPaths(); ++i) {
      PathPtr_t p = pv->pathAtRank(i);
      if (PathVectorPtr_t sub = HPP_DYNAMIC_PTR_CAST(PathVector, p)) { if (!flatten(sub, out)) return false; continue; }
      StraightPathPtr_t sp = HPP_DYNAMIC_PTR_CAST(StraightPath, p);
      if (!sp || sp->constraints()) return false;
      out.push_back(sp);
    }
    return true;
  }
  ValidationReportPtr_t configReport(int reason, int pair, ConfigurationIn_t qFail) {
    if (reason == HPPB200_REASON_JOINT_BOUND) return jointBoundReport(scene_->robot(), pair, qFail[pair]);
    hppb200_model* m = scene_->model();
    const bool disc = method_ == HPPB200_METHOD_DISCRETIZED;
    std::vector<int32_t> A(disc ? hppb200_model_num_config_pairs(m) : hppb200_model_num_body_geom_pairs(m)), B(A.size());
    if (disc) hppb200_model_get_config_pairs(m, A.data(), B.data(), nullptr);
    else hppb200_model_get_body_geom_pairs(m, A.data(), B.data(), nullptr);
    return collisionReport(*scene_, A, B, reason == HPPB200_REASON_COLLISION ? pair : -1);
  }
  bool validateStraight(const StraightPathPtr_t& sp, bool reverse, PathPtr_t& validPart, PathValidationReportPtr_t& report) {
    const interval_t tr = sp->timeRange();
    const Configuration_t q1 = sp->initial(), q2 = sp->end();
    const double L = tr.second - tr.first;
    uint8_t valid, reason; int32_t pair; double tFail, tLast;
    {
      Scene::Pool::Lease ws = scene_->workspace();
      check(hppb200_validate_edges(ws.get(), method_, param_, reverse ? 1 : 0, q1.data(), q2.data(), &L, 1, 0, &valid, &tFail, &tLast,
                                   &pair, &reason));
    }
    if (valid) { validPart = sp; return true; }
    if (reason == HPPB200_REASON_ITER_CAP)  // the engine's iteration cap (the reference has none): let the stock validator finish it
      return stock_->validate(sp, reverse, validPart, report);
    // ownership rule: validPart is set on every return (progressive.cc:121-127, discretized.cc:99-105)
    validPart = reverse ? sp->extract(tr.first + tLast, tr.second) : sp->extract(tr.first, tr.first + tLast);
    bool ok = true;
    report = PathValidationReportPtr_t(new PathValidationReport(tr.first + tFail, configReport(reason, pair, sp->eval(tr.first + tFail, ok))));
    return false;
  }
  // path::Spline<BernsteinBasis, 3>: base() [nq], parameters() [4][nv] (Eigen row-major matrix, rows = control points)
  bool validateSpline(const shared_ptr<path::Spline<path::BernsteinBasis, 3> >& sp, bool reverse, PathPtr_t& validPart,
                      PathValidationReportPtr_t& report) {
    const interval_t tr = sp->timeRange();
    const Configuration_t base = sp->base();
    const Eigen::Matrix<value_type, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> P = sp->parameters();
    const double L = tr.second - tr.first;
    uint8_t valid, reason; int32_t pair; double tFail, tLast;
    {
      Scene::Pool::Lease ws = scene_->workspace();
      check(hppb200_validate_spline_edges(ws.get(), method_, param_, reverse ? 1 : 0, base.data(), P.data(), &L, 1, 0, &valid, &tFail, &tLast,
                                          &pair, &reason));
    }
    if (valid) { validPart = sp; return true; }
    if (reason == HPPB200_REASON_ITER_CAP) return stock_->validate(sp, reverse, validPart, report);
    validPart = reverse ? sp->extract(tr.first + tLast, tr.second) : sp->extract(tr.first, tr.first + tLast);
    bool ok = true;
    report = PathValidationReportPtr_t(new PathValidationReport(tr.first + tFail, configReport(reason, pair, sp->eval(tr.first + tFail, ok))));
    return false;
  }
  // src/continuous-validation.cc:153-192 with every leaf validated in one batch
  bool validateChain(const PathVectorPtr_t& pv, const std::vector<StraightPathPtr_t>& leaves, bool reverse, PathPtr_t& validPart,
                     PathValidationReportPtr_t& report) {
    const std::size_t n = leaves.size(), nq = (std::size_t)pv->outputSize();
    std::vector<double> q1(n * nq), q2(n * nq), L(n);
    for (std::size_t i = 0; i < n; ++i) {
      const Configuration_t a = leaves[i]->initial(), b = leaves[i]->end();
      std::copy(a.data(), a.data() + nq, q1.begin() + i * nq);
      std::copy(b.data(), b.data() + nq, q2.begin() + i * nq);
      L[i] = leaves[i]->length();
    }
    const int32_t chain[2] = {0, (int32_t)n};
    uint8_t valid, reason; int32_t pair, bad; double tFail, tLast;
    {
      Scene::Pool::Lease ws = scene_->workspace();
      check(hppb200_validate_edge_chains(ws.get(), method_, param_, reverse ? 1 : 0, q1.data(), q2.data(), L.data(), n, chain, 1, 0, &valid,
                                         &tFail, &tLast, &pair, &reason, &bad));
    }
    PathVectorPtr_t out = PathVector::create(pv->outputSize(), pv->outputDerivativeSize());
    validPart = out;
    if (valid) {  // the reference returns a PathVector of copies (:163,183), not `path` itself
      for (const StraightPathPtr_t& l : leaves) out->appendPath(l->copy());
      return true;
    }
    if (reason == HPPB200_REASON_ITER_CAP) return stock_->validate(pv, reverse, validPart, report);
    // the failing leaf once more, alone: its own valid part and report in its own parameterisation
    PathPtr_t localValid;
    PathValidationReportPtr_t localReport;
    validateStraight(leaves[(std::size_t)bad], reverse, localValid, localReport);
    double before = 0;
    for (int i = 0; i < bad; ++i) before += L[(std::size_t)i];
    localReport->parameter = tFail;  // = local parameter + offset, accumulated on the device like the reference does (:167,186)
    report = localReport;
    if (!reverse) {
      for (int i = 0; i < bad; ++i) out->appendPath(leaves[(std::size_t)i]->copy());
      out->appendPath(localValid->copy());
    } else {
      out->appendPath(localValid->copy());
      for (std::size_t i = (std::size_t)bad + 1; i < n; ++i) out->appendPath(leaves[i]->copy());
    }
    (void)before; (void)tLast;
    return false;
  }
  ScenePtr_t scene_;
  int method_;
  value_type param_;
  PathValidationPtr_t stock_;
};

// ---- NearestNeighbor --------------------------------------------------------------------------------
// Replaces nearestNeighbor::Basic (src/nearest-neighbor/basic.cc:52-181): nodes are mirrored on the device in
// insertion order, connected components as integer labels DERIVED from the roadmap before every component query
// (ComponentLabels in b200_adapter_util.hh: the reference never calls NearestNeighbor::merge, src/roadmap.cc:298-310).
// The metric is the problem's WeighedDistance: its weights() go into the engine's model (setDistance); with any other
// distance type the plugin keeps the stock nearest-neighbour object (see the planner wrappers below).
class B200NearestNeighbor : public NearestNeighbor {
  typedef ComponentLabels<ConnectedComponent, ConnectedComponentPtr_t, ConnectedComponentWkPtr_t, Node> Labels_t;

 public:
  B200NearestNeighbor(const DistancePtr_t& d, const ScenePtr_t& scene) : distance_(d), scene_(scene) { init(); }
  ~B200NearestNeighbor() override { if (nn_) hppb200_knn_destroy(nn_); }
  void clear() override {
    if (nn_) hppb200_knn_destroy(nn_);
    nn_ = nullptr;
    nodes_.clear();
    if (labels_) labels_->clear();
    init();
  }
  void addNode(const NodePtr_t& node) override {
    ready();
    const int32_t cc = labels_->addNode(node, node->connectedComponent());
    check(hppb200_knn_append(nn_, node->configuration().data(), &cc, 1));
    nodes_.push_back(node);
  }
  NodePtr_t search(ConfigurationIn_t q, const ConnectedComponentPtr_t& cc, value_type& d, bool) override {
    ready();
    int64_t idx; check(hppb200_knn_query(nn_, q.data(), 1, 1, labels_->sync(cc), 0, &idx, &d));
    return idx < 0 ? NodePtr_t() : nodes_[(std::size_t)idx];  // WeighedDistance is symmetric: `reverse` is a no-op
  }
  NodePtr_t search(const NodePtr_t& n, const ConnectedComponentPtr_t& cc, value_type& d) override {
    return search(n->configuration(), cc, d, false);
  }
  Nodes_t KnearestSearch(ConfigurationIn_t q, const ConnectedComponentPtr_t& cc, const std::size_t K, value_type& d) override {
    ready();
    return knn(q, labels_->sync(cc), K, d);
  }
  Nodes_t KnearestSearch(const NodePtr_t& n, const ConnectedComponentPtr_t& cc, const std::size_t K, value_type& d) override {
    ready();
    return knn(n->configuration(), labels_->sync(cc), K, d);
  }
  Nodes_t KnearestSearch(ConfigurationIn_t q, const RoadmapPtr_t& roadmap, const std::size_t K, value_type& d) override {
    ready();
    if (roadmap && roadmap->nodes().size() != nodes_.size())  // a restored roadmap: adopt its nodes component by component
      for (const ConnectedComponentPtr_t& cc : roadmap->connectedComponents()) labels_->sync(cc);
    return knn(q, -1, K, d);
  }
  NodeVector_t withinBall(ConfigurationIn_t q, const ConnectedComponentPtr_t& cc, value_type maxD) override {
    ready();
    const int32_t l = labels_->sync(cc);
    std::vector<int64_t> idx(nodes_.size()); int64_t count = 0;
    check(hppb200_knn_within_ball(nn_, q.data(), 1, maxD, l, idx.size(), idx.data(), &count));
    NodeVector_t out;
    for (int64_t i = 0; i < count; ++i) out.push_back(nodes_[(std::size_t)idx[i]]);
    return out;
  }
  // never called by Roadmap (Basic::merge is empty too, src/nearest-neighbor/basic.hh:81); labels follow the roadmap anyway
  void merge(ConnectedComponentPtr_t cc1, ConnectedComponentPtr_t) override { ready(); labels_->sync(cc1); }
  DistancePtr_t distance() const override { return distance_; }

  // batch extension (kPRM*-style linking): nearest first, -1 padded; see hppb200_knn_query
  void knnBatch(const double* queries, std::size_t n, int K, const ConnectedComponentPtr_t& cc, int64_t* idx, double* dist) {
    ready();
    check(hppb200_knn_query(nn_, queries, n, K, cc ? labels_->sync(cc) : -1, 0, idx, dist));
  }

 private:
  B200NearestNeighbor() {}  // serialization only: the mirror is rebuilt from the restored roadmap at the first query
  HPP_SERIALIZABLE();
  void init() {
    if (!scene_) {
      WeighedDistancePtr_t wd = HPP_DYNAMIC_PTR_CAST(WeighedDistance, distance_);
      if (!wd) throw std::runtime_error("B200NearestNeighbor needs a WeighedDistance");
      scene_.reset(new Scene(wd->robot()));
    }
    if (!scene_->setDistance(distance_)) throw std::runtime_error("B200NearestNeighbor needs a WeighedDistance over the scene's robot");
    check(hppb200_knn_create(scene_->model(), 0, &nn_));
    labels_.reset(new Labels_t([this](int32_t from, int32_t to) { check(hppb200_knn_relabel(nn_, from, to)); },
                               [this](Node* n, int32_t l) {
                                 check(hppb200_knn_append(nn_, n->configuration().data(), &l, 1));
                                 nodes_.push_back(n);
                               }));
  }
  void ready() { if (!nn_) init(); }
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
  std::vector<NodePtr_t> nodes_;  // mirror index -> node (NodePtr_t is a raw Node*, owned by the roadmap)
  std::unique_ptr<Labels_t> labels_;
};

// like nearestNeighbor::Basic (src/nearest-neighbor/serialization.cc:24-33): the distance only; nodes come back
// through the roadmap (src/serialization.cc:86-99)
template <typename Archive>
inline void B200NearestNeighbor::serialize(Archive& ar, const unsigned int version) {
  (void)version;
  ar& boost::serialization::make_nvp("base", boost::serialization::base_object<NearestNeighbor>(*this));
  ar& boost::serialization::make_nvp("distance_", distance_);
}
HPP_SERIALIZATION_IMPLEMENT(B200NearestNeighbor);

// ---- registration (reference src/plugin.cc:43-104, include/hpp/core/problem-solver.hh:52-59,519-522) ----
class B200Plugin : public ProblemSolverPlugin {
 public:
  B200Plugin() : ProblemSolverPlugin("B200ValidationPlugin", "0.2") {}

 protected:
  bool impl_initialize(ProblemSolverPtr_t ps) override {
    ps->configValidations.add("B200CollisionValidation", [](const DevicePtr_t& r) -> ConfigValidationPtr_t { return B200ConfigValidation::create(r, false); });
    ps->configValidations.add("B200ConfigValidations", [](const DevicePtr_t& r) -> ConfigValidationPtr_t { return B200ConfigValidation::create(r, true); });
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
        // Roadmap::nearestNeighbor requires an empty roadmap (src/roadmap.cc:74-84).  Only a WeighedDistance is the
        // engine's metric: with any other distance the stock nearestNeighbor::Basic stays.
        if (roadmap->nodes().empty() && HPP_DYNAMIC_PTR_CAST(WeighedDistance, problem->distance())) {
          ScenePtr_t scene;
          if (shared_ptr<B200PathValidation> pv = HPP_DYNAMIC_PTR_CAST(B200PathValidation, problem->pathValidation())) scene = pv->scene();
          else scene.reset(new Scene(problem->robot()));
          if (scene->setDistance(problem->distance())) roadmap->nearestNeighbor(new B200NearestNeighbor(problem->distance(), scene));
        }
        return stock(problem, roadmap);
      });
    }
    return true;
  }
};
}  // namespace b200
}  // namespace hpp

BOOST_CLASS_EXPORT(hpp::b200::B200NearestNeighbor)
HPP_CORE_DEFINE_PLUGIN(hpp::b200::B200Plugin)
