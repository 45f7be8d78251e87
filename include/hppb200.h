/*
 * hppb200.h — C ABI of the B200-native batched validation engine.
 *
 * This is the drop-in boundary for hpp-core's data-parallel hot path
 * (SURVEY.md §8b).  Every entry point takes plain pointers and sizes, returns
 * a status code (no exceptions cross the ABI) and is what a hpp-core adapter
 *  binds in place of the reference virtuals:
 *
 *   hppb200_validate_configs   <- hpp::core::ConfigValidations::validate /
 *                                 CollisionValidation::validate /
 *                                 JointBoundValidation::validate
 *                                 (reference src/config-validations.cc:40-49,
 *                                  src/collision-validation.cc:48-95,
 *                                  src/joint-bound-validation.cc:45-80)
 *   hppb200_validate_edges     <- PathValidation::validate for
 *                                 pathValidation::Discretized
 *                                 (src/path-validation/discretized.cc:55-107),
 *                                 continuousValidation::Progressive
 *                                 (src/continuous-validation/progressive.cc:78-129),
 *                                 continuousValidation::Dichotomy
 *                                 (src/continuous-validation/dichotomy.cc:68-132)
 *   hppb200_straight_lengths   <- steeringMethod::Straight::impl_compute
 *                                 (src/steering-method/straight.cc:38-51) /
 *                                 WeighedDistance::impl_distance
 *                                 (src/weighed-distance.cc:272-289)
 *   hppb200_path_eval          <- Path::eval (include/hpp/core/path.hh:135-140) ->
 *                                 StraightPath::impl_compute (src/straight-path.cc:91-110),
 *                                 which StraightPath::impl_extract (:142-158) is built on
 *   hppb200_knn_*              <- nearestNeighbor::Basic::{search,KnearestSearch,
 *                                 withinBall} (src/nearest-neighbor/basic.cc:52-181)
 *   hppb200_model_*            <- ObstacleUser::{addRobotCollisionPairs,addObstacle,
 *                                 setSecurityMargins} (src/obstacle-user.cc:53-241),
 *                                 ContinuousValidation::{Initialize,AddObstacle}
 *                                 (src/continuous-validation.cc:63-118),
 *                                 SolidSolidCollision ctor
 *                                 (src/continuous-validation/solid-solid-collision.cc:162-299)
 *
 * All floating-point data is IEEE-754 binary64 (hpp::core::value_type,
 * reference include/hpp/core/fwd.hh:174).  Matrices are row-major.
 * A rigid placement is 12 doubles: R (9, row-major) then t (3).
 *
 * There is no CPU fallback behind this ABI: every compute entry point needs a
 * CUDA device and fails with HPPB200_ERR_CUDA otherwise.
 */
#ifndef HPPB200_H
#define HPPB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HPPB200_ABI_VERSION 2

/* ---- status codes ------------------------------------------------------ */
enum {
  HPPB200_OK = 0,
  HPPB200_ERR_INVALID_ARGUMENT = 1, /* std::invalid_argument in the reference */
  HPPB200_ERR_RUNTIME = 2,          /* std::runtime_error in the reference    */
  HPPB200_ERR_CUDA = 3,             /* no device / CUDA failure               */
  HPPB200_ERR_UNSUPPORTED = 4
};

/* ---- joint types (pinocchio joint models reachable from hpp-pinocchio) -- */
enum {
  HPPB200_JOINT_REVOLUTE_X = 0, /* bounded revolute, nq = nv = 1 (R^1, rotation) */
  HPPB200_JOINT_REVOLUTE_Y = 1,
  HPPB200_JOINT_REVOLUTE_Z = 2,
  HPPB200_JOINT_PRISMATIC_X = 3, /* nq = nv = 1 (R^1, translation) */
  HPPB200_JOINT_PRISMATIC_Y = 4,
  HPPB200_JOINT_PRISMATIC_Z = 5,
  HPPB200_JOINT_FREEFLYER = 6 /* R^3 x SO(3): nq = 7 (x y z qx qy qz qw, unit quaternion), nv = 6 (dt, log3) */
};

/* ---- shape types (coal::ShapeBase subclasses used by the path) ---------- */
enum {
  HPPB200_SHAPE_SPHERE = 0,  /* params: radius                                 */
  HPPB200_SHAPE_CAPSULE = 1, /* params: radius, halfLength (axis = local z)    */
  HPPB200_SHAPE_BOX = 2,     /* params: half extents hx, hy, hz                */
  HPPB200_SHAPE_CONVEX = 3   /* vertices [vert_begin, vert_begin+vert_count)   */
};

/* ---- edge validation methods ------------------------------------------- */
enum {
  HPPB200_METHOD_DISCRETIZED = 0, /* param = step size (> 0)  */
  HPPB200_METHOD_PROGRESSIVE = 1, /* param = tolerance (> 0)  */
  HPPB200_METHOD_DICHOTOMY = 2    /* param = tolerance (== 0) */
};

/* ---- per-item result codes written to `status`/`reason` arrays ---------- */
enum {
  HPPB200_REASON_VALID = 0,
  HPPB200_REASON_COLLISION = 1,   /* CollisionValidationReport        */
  HPPB200_REASON_JOINT_BOUND = 2, /* JointBoundValidationReport       */
  HPPB200_REASON_ITER_CAP = 3     /* continuous validation hit max_iterations
                                     (the reference has no cap; flagged, never
                                     silently accepted)                        */
};

/*
 * Robot + environment description, the flattened form of what hpp-core reads
 * from pinocchio::Model / GeomModel (robot) and from the obstacle GeomModel
 * (ProblemSolver::addObstacle, reference src/problem-solver.cc:863-901).
 * Joint indices are pinocchio's: 0 is the universe, moving joints are
 * 1..njoints and parent[i] < i.  Arrays indexed by joint have njoints entries,
 * entry k describing joint k+1.
 */
typedef struct hppb200_model_desc {
  int32_t njoints;
  const int32_t* joint_parent;    /* [njoints] parent joint index (0 = universe)   */
  const int32_t* joint_type;      /* [njoints] HPPB200_JOINT_*                     */
  const double* joint_placement;  /* [njoints][12] placement in the parent frame   */
  const double* q_lower;          /* [nq] lower configuration bounds               */
  const double* q_upper;          /* [nq] upper configuration bounds               */

  int32_t ngeoms;                 /* robot collision geometries (pinocchio order)  */
  const int32_t* geom_joint;      /* [ngeoms] parent joint (0 = fixed to universe) */
  const double* geom_placement;   /* [ngeoms][12] placement in the joint frame     */
  const int32_t* geom_shape;      /* [ngeoms] HPPB200_SHAPE_*                      */
  const double* geom_params;      /* [ngeoms][4]                                   */
  const int32_t* geom_vert_begin; /* [ngeoms] (convex only, else 0)                */
  const int32_t* geom_vert_count; /* [ngeoms]                                      */

  int32_t nself_pairs;            /* active GeomModel::collisionPairs, in order    */
  const int32_t* self_pairs;      /* [nself_pairs][2] geometry indices             */

  int32_t nobstacles;             /* obstacles in addObstacle() call order         */
  const double* obst_placement;   /* [nobstacles][12] world placement              */
  const int32_t* obst_shape;      /* [nobstacles]                                  */
  const double* obst_params;      /* [nobstacles][4]                               */
  const int32_t* obst_vert_begin; /* [nobstacles]                                  */
  const int32_t* obst_vert_count; /* [nobstacles]                                  */

  int32_t nvertices;
  const double* vertices;         /* [nvertices][3] shared convex vertex pool      */

  /* (njoints+1)x(njoints+1) security-margin matrix indexed by joint index
     (ObstacleUser::setSecurityMargins, reference src/obstacle-user.cc:166-196);
     NULL = all zero.  Obstacles use row/column 0. */
  const double* security_margins;

  double break_distance;          /* ContinuousValidation::breakDistance, default 1e-2
                                     (reference src/continuous-validation.cc:411);
                                     only affects BVH pairs in coal, kept for API
                                     fidelity */
  const double* weights;          /* [njoints] WeighedDistance::weights(); NULL = 1 */

  /* ---- pair-table edits (ABI version 2; all zero = none) -------------------------------
     obst_global[k] != 0 (or obst_global == NULL): obstacle k was passed to addObstacle()
     (paired with every joint, reference src/obstacle-user.cc:53-58, continuous-validation.cc
     :103-118); 0: the object only exists for HPPB200_OP_ADD_OBSTACLE_TO_JOINT.
     pair_ops are applied in order AFTER the base tables (self pairs, global obstacles,
     security_margins) are built; the adapter keeps the log of the calls it received through
     ObstacleUserInterface and rebuilds the model when it changes. */
  const uint8_t* obst_global;     /* [nobstacles] or NULL                           */
  int32_t npair_ops;
  const struct hppb200_pair_op* pair_ops;
  int32_t check_parameterized;    /* CollisionValidation::checkParameterized_       */
} hppb200_model_desc;

/* One recorded call of the ObstacleUserInterface (reference include/hpp/core/obstacle-user.hh:47-88,
   src/obstacle-user.cc:60-220, src/continuous-validation.cc:224-367).  Geometry references: >= 0 robot
   geometry, -(k+1) obstacle k.  Joint indices are pinocchio's (1..njoints). */
#define HPPB200_OP_ADD_OBSTACLE_TO_JOINT 1      /* a = obstacle k, b = joint, flag = includeChildren; configuration
                                                   validation only (ContinuousValidation has no such call)        */
#define HPPB200_OP_REMOVE_OBSTACLE_FROM_JOINT 2 /* a = obstacle k, b = joint; not registered => HPPB200_ERR_RUNTIME */
#define HPPB200_OP_FILTER_COLLISION_PAIRS 3     /* data = int32 [(njoints+1)^2] row-major RelativeMotion matrix:
                                                   0 Constrained (disabled), 1 Parameterized, 2 Unconstrained    */
#define HPPB200_OP_SET_SECURITY_MARGINS 4       /* data = double [(njoints+1)^2] row-major                       */
#define HPPB200_OP_SET_MARGIN_BETWEEN_BODIES 5  /* a, b = geometry references, value = margin; configuration table:
                                                   first matching pair, none => HPPB200_ERR_INVALID_ARGUMENT;
                                                   continuous table: every matching pair                         */
#define HPPB200_OP_ADD_OBSTACLE 6               /* a = obstacle k (obst_global[k] = 0): addObstacle() called at this point
                                                   of the log, i.e. after earlier filters / margin matrices, which
                                                   therefore do not apply to its pairs (obstacle-user.cc:53-58,
                                                   continuous-validation.cc:103-118)                              */
typedef struct hppb200_pair_op {
  int32_t kind, a, b, flag;
  double value;
  const void* data;
} hppb200_pair_op;

typedef struct hppb200_model hppb200_model;         /* immutable compiled tables   */
typedef struct hppb200_workspace hppb200_workspace; /* one CUDA stream + scratch   */
typedef struct hppb200_knn hppb200_knn;             /* device mirror of roadmap nodes */

typedef struct hppb200_stats {
  uint64_t configs;          /* configurations run through FK                  */
  uint64_t pair_tests;       /* narrow-phase calls actually executed           */
  uint64_t gjk_iterations;   /* summed GJK iterations                          */
  uint64_t edge_samples;     /* path samples evaluated by edge validation      */
  uint64_t kernel_launches;  /* CUDA kernels launched by this workspace        */
  uint64_t iter_cap_hits;    /* edges that ended with HPPB200_REASON_ITER_CAP  */
  double last_kernel_ms;     /* device time of the last validate_* call        */
  uint64_t engine_ticks;     /* continuous engine: control/narrow-phase rounds  */
  uint64_t discarded_tests;  /* speculative pair tests whose result the reference
                                order did not need (counted in pair_tests too)  */
} hppb200_stats;

/* ---- library ----------------------------------------------------------- */
int hppb200_abi_version(void);
const char* hppb200_last_error(void); /* thread-local message of the last failure */
int hppb200_device_count(void);       /* 0 when no CUDA device is visible        */
/* Measured FP64 FMA throughput of `device` in TFLOP/s (independent DFMA chains, CUDA events):
   the roofline denominator for the narrow-phase kernels. */
int hppb200_measure_fp64_peak(int device, double* tflops);

/* Measured TMEM -> register bandwidth of `device` in TB/s (tcgen05.ld.32x32b.x32 from 16 warps per SM over a resident
   512-column allocation, CUDA events): the roofline denominator of the k-NN tensor-core filter pass, whose every
   (query, node) accumulator word has to leave tensor memory that way. */
int hppb200_measure_tmem_read_peak(int device, double* tb_per_s);

/* ---- model ------------------------------------------------------------- */
int hppb200_model_create(const hppb200_model_desc* desc, hppb200_model** out);
void hppb200_model_destroy(hppb200_model* m);
int hppb200_model_nq(const hppb200_model* m);
int hppb200_model_nv(const hppb200_model* m);
/* Configuration-validation pair table in cPairs_ order:
   geomA is a robot geometry, geomB >= 0 a robot geometry or -(k+1) for obstacle k. */
int hppb200_model_num_config_pairs(const hppb200_model* m);
int hppb200_model_get_config_pairs(const hppb200_model* m, int32_t* geomA,
                                   int32_t* geomB, double* margin);
/* How many trailing entries of that table are pPairs_ (parameterized pairs, tested after cPairs_ when
   check_parameterized is set; reference src/collision-validation.cc:62-67); 0 otherwise. */
int hppb200_model_num_parameterized_pairs(const hppb200_model* m);
/* Continuous-validation body pairs in intervalValidations_ order. */
int hppb200_model_num_body_pairs(const hppb200_model* m);
/* joint_a/joint_b [nbp]; geom-pair range [pair_begin[i], pair_begin[i+1]) into
   (pairA, pairB); coefficient range [coef_begin[i], coef_begin[i+1]) a->b and
   [rcoef_begin[i], rcoef_begin[i+1]) b->a.  Any output pointer may be NULL. */
int hppb200_model_get_body_pairs(const hppb200_model* m, int32_t* joint_a,
                                 int32_t* joint_b, int32_t* pair_begin,
                                 int32_t* coef_begin, int32_t* rcoef_begin);
int hppb200_model_num_body_geom_pairs(const hppb200_model* m);
int hppb200_model_get_body_geom_pairs(const hppb200_model* m, int32_t* geomA,
                                      int32_t* geomB, double* margin);
int hppb200_model_num_coefficients(const hppb200_model* m, int reverse);
int hppb200_model_get_coefficients(const hppb200_model* m, int reverse,
                                   int32_t* joint, double* value);
/* Body::radius() per joint index 0..njoints (pinocchio GeomData::radius). */
int hppb200_model_get_body_radii(const hppb200_model* m, double* radius);

/* ---- workspace --------------------------------------------------------- */
int hppb200_workspace_create(const hppb200_model* m, int device,
                             hppb200_workspace** out);
void hppb200_workspace_destroy(hppb200_workspace* ws);
int hppb200_get_stats(const hppb200_workspace* ws, hppb200_stats* out);
/* Use an externally owned cudaStream_t (e.g. torch's current stream). */
int hppb200_workspace_set_stream(hppb200_workspace* ws, void* cuda_stream);

/* ---- configuration validation ----------------------------------------- */
#define HPPB200_CHECK_JOINT_BOUNDS 1u /* JointBoundValidation first, as in the default
                                         Problem (reference src/problem.cc)          */
#define HPPB200_DEVICE_POINTERS 2u    /* in/out arrays already live on the device    */

/* q [n][nq]; valid [n] (1 valid / 0 invalid); first_pair [n] index of the first
   colliding pair in cPairs_ order, or the violated configuration rank for a joint
   bound, -1 when valid; reason [n] HPPB200_REASON_* (may be NULL). */
int hppb200_validate_configs(hppb200_workspace* ws, const double* q, size_t n,
                             uint32_t flags, uint8_t* valid, int32_t* first_pair,
                             uint8_t* reason);

/* Same result as hppb200_validate_configs, for callers that cannot allocate device memory (CUDA graph capture, serving
   loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
   Calls no cudaMalloc / cudaFree / allocating library routine: the kernels' scratch (job counter, job
   queue, geometry placements) lives in ONE static device array of HPPB200_NOALLOC_SCRATCH_BYTES per GPU, held by every
   GPU that loads the library until the process exits.  Requirements: flags must contain HPPB200_DEVICE_POINTERS (q,
   valid, first_pair, reason all device pointers, all non-NULL); n limited by the array:
   256 + 8 (2 n + 4096) + 96 (ngeoms [+ njoints for tree-shaped robots]) n bytes must fit, i.e. about 85 000
   configurations per call for UR5 (8 geometries); more => HPPB200_ERR_INVALID_ARGUMENT with the limit in the message.
   The array is shared by every workspace and call on that GPU and keeps the last call's contents (the call clears the
   job counter itself, nothing else needs to start at zero): calls that overlap on the GPU (two workspaces / streams /
   threads) race on it - serialise them.  The first workspace created on a device resolves the array's address. */
#define HPPB200_NOALLOC_SCRATCH_BYTES (64u << 20)
int hppb200_validate_configs_noalloc(hppb200_workspace* ws, const double* q, size_t n,
                                     uint32_t flags, uint8_t* valid, int32_t* first_pair,
                                     uint8_t* reason);

/* Forward kinematics only: oMg [n][ngeoms][12] robot geometry world placements
   (pinocchio::forwardKinematics + updateGeometryPlacements). Test/diagnostic. */
int hppb200_geometry_placements(hppb200_workspace* ws, const double* q, size_t n,
                                uint32_t flags, double* oMg);

/* Signed distance minus security margin (coal distance_lower_bound semantics for
   shape pairs) of every configuration-validation pair: dist [n][npairs].
   Test/diagnostic; mirrors DistanceBetweenObjects (src/distance-between-objects.cc). */
int hppb200_pair_distances(hppb200_workspace* ws, const double* q, size_t n,
                           uint32_t flags, double* dist);

/* CollisionValidation with computeAllContacts (reference src/collision-validation.cc:72-89,
   include/hpp/core/collision-validation.hh:63-65): what AllCollisionsValidationReport::collisionReports lists, per
   configuration: mask [n][(npairs + 31) / 32], bit p set iff pair p of the configuration-pair table
   (hppb200_model_get_config_pairs) collides AND lies in the table (cPairs_ / pPairs_) of the first colliding pair, at
   or after it; count [n] = number of set bits; valid / first_pair as hppb200_validate_configs (collision only). */
int hppb200_all_collisions(hppb200_workspace* ws, const double* q, size_t n, uint32_t flags,
                           uint8_t* valid, int32_t* first_pair, uint32_t* mask, int32_t* count);

/* DistanceBetweenObjects::computeDistances (reference src/distance-between-objects.cc:41-77): for every configuration the
   distance, without security margin, between every inner object of every moving body and every obstacle:
   dist [n][npairs], pairs obstacle-major in addObstacle order, then joints, then the joint's inner objects
   (hppb200_model_num_distance_pairs of them).  Separated shapes: the exact separation (what
   coal::DistanceResult::min_distance holds); overlapping shapes: core distance 0 minus the inflation radii (no EPA
   penetration depth). */
int hppb200_model_num_distance_pairs(const hppb200_model* m);
int hppb200_distance_between_objects(hppb200_workspace* ws, const double* q, size_t n, uint32_t flags, double* dist);

/* ---- edge (StraightPath) validation ------------------------------------ */
/* L[i] = WeighedDistance(q1[i], q2[i]) (steeringMethod::Straight). */
int hppb200_straight_lengths(hppb200_workspace* ws, const double* q1,
                             const double* q2, size_t n, uint32_t flags, double* L);

/* q[i] = configuration of edge i at parameter t[i] in [0, L[i]]: the stored end points bit-exactly at
   t = 0 / t = L (and for L = 0), LiegroupSpace::interpolate(q1, q2, t / L) otherwise (R^n joints linear,
   free-flyer R^3 x SO(3)).  The adapter builds validPart = extract(0, t_last_valid) from it. */
int hppb200_path_eval(hppb200_workspace* ws, const double* q1, const double* q2, const double* L,
                      const double* t, size_t n, uint32_t flags, double* q);

/* Edge i is the StraightPath from q1[i] to q2[i] over [0, L[i]].
   valid [n]; t_fail [n] report->parameter (NaN when valid); t_last_valid [n] end of
   the valid part (forward: validPart = extract(0, t_last_valid); reverse:
   extract(t_last_valid, L)); pair [n] colliding pair (cPairs_ index for Discretized,
   body-pair geometry-pair index for continuous methods; configuration rank for a
   joint bound), -1 when valid; reason [n] HPPB200_REASON_*. */
int hppb200_validate_edges(hppb200_workspace* ws, int method, double param,
                           int reverse, const double* q1, const double* q2,
                           const double* L, size_t n, uint32_t flags,
                           uint8_t* valid, double* t_fail, double* t_last_valid,
                           int32_t* pair, uint8_t* reason);
/* Edges that are path::Spline<BernsteinBasis, 3> (reference include/hpp/core/path/spline.hh, src/path/spline.cc; what
   steeringMethod::Spline<BernsteinBasis, 3>::steer returns, src/steering-method/spline.cc:75-142, and the path type of
   the reference's non-straight validation test, tests/test-continuous-validation.cc:288-427): spline i is
   q(t) = base_i + sum_k b_k((t - 0) / L_i) params_i[k], t in [0, L_i], b_k the cubic Bernstein polynomials; base
   [n][nq] = Spline::base(), params [n][4][nv] row-major = Spline::parameters(), L [n] = the time range.  Same methods,
   outputs and error behaviour as hppb200_validate_edges.  The continuous methods use Spline::impl_velocityBound on
   sub-intervals with the two-round refinement of BodyPairCollision::collisionFreeInterval (refine_ = true,
   src/continuous-validation/body-pair-collision.cc:101-173).  Vector-space configuration spaces only (nq == nv: no
   free-flyer joint), else HPPB200_ERR_INVALID_ARGUMENT. */
int hppb200_validate_spline_edges(hppb200_workspace* ws, int method, double param, int reverse, const double* base,
                                  const double* params, const double* L, size_t n, uint32_t flags, uint8_t* valid,
                                  double* t_fail, double* t_last_valid, int32_t* pair, uint8_t* reason);
/* A PathVector of StraightPaths per chain (what RandomShortcut and the other optimisers hand to the path validation):
   chain c = edges [chain_begin[c], chain_begin[c+1]) in path order; chain_begin has n_chains + 1 entries.  Replays the
   recursion of ContinuousValidation::validate (reference src/continuous-validation.cc:153-192) for Progressive and
   Dichotomy: all leaves are validated as one batch, then per chain the first invalid leaf in walking order (last to
   first when reverse) decides.  Outputs per chain, in the PathVector's own parameterisation: valid; t_fail =
   report->parameter (NaN when valid); t_last_valid = end of the valid part (forward: validPart = extract(0,
   t_last_valid); reverse: start of the valid part); pair, reason of the failing leaf; first_bad_edge = its edge index
   (-1 when valid).  Discretized treats a PathVector as one path: HPPB200_ERR_RUNTIME. */
int hppb200_validate_edge_chains(hppb200_workspace* ws, int method, double param, int reverse,
                                 const double* q1, const double* q2, const double* L, size_t n_edges,
                                 const int32_t* chain_begin, size_t n_chains, uint32_t flags,
                                 uint8_t* valid, double* t_fail, double* t_last_valid, int32_t* pair,
                                 uint8_t* reason, int32_t* first_bad_edge);
/* Iteration cap per edge for continuous methods (default 1<<20). */
int hppb200_workspace_set_max_iterations(hppb200_workspace* ws, uint32_t max_iter);

/* ---- multi-GPU: a device list, sharded internally ------------------------ */
/* Configurations and edges are independent units (the reference validates them from an OpenMP loop with pooled state,
   tests/test-continuous-validation.cc:324-343; src/continuous-validation.cc:193-205): a handle owns one workspace and
   one persistent host thread per device; every call splits the caller's HOST arrays into contiguous index ranges
   [r n / N, (r + 1) n / N), validates them concurrently through the single-GPU host-pointer path and writes the results
   into the caller's output arrays in place (that placement is the gather of the validity masks; no collective).
   Same outputs as the single-GPU entry points.  devices == NULL / ndevices == 0: every visible device.
   HPPB200_DEVICE_POINTERS is rejected.  One call at a time per handle (calls are serialised internally). */
typedef struct hppb200_multi hppb200_multi;
int hppb200_multi_create(const hppb200_model* m, const int* devices, int ndevices, hppb200_multi** out);
void hppb200_multi_destroy(hppb200_multi* h);
int hppb200_multi_num_devices(const hppb200_multi* h);
/* borrowed: the workspace of shard `rank` (statistics, iteration cap); NULL if out of range */
hppb200_workspace* hppb200_multi_workspace(hppb200_multi* h, int rank);
int hppb200_multi_validate_configs(hppb200_multi* h, const double* q, size_t n, uint32_t flags,
                                   uint8_t* valid, int32_t* first_pair, uint8_t* reason);
int hppb200_multi_validate_edges(hppb200_multi* h, int method, double param, int reverse,
                                 const double* q1, const double* q2, const double* L, size_t n,
                                 uint32_t flags, uint8_t* valid, double* t_fail, double* t_last_valid,
                                 int32_t* pair, uint8_t* reason);

/* Path validation of edges between ROADMAP NODES (kPRM*-style linking, reference src/path-planner/k-prm-star.cc:134-193:
   (*sm)(q_i, q_j) then pathValidation()->validate for nodes i, j that are already in the roadmap): node_a / node_b [n] are
   indices into the device mirror `nn` (insertion order, as hppb200_knn_query returns them); the end configurations are
   gathered on the device and L = WeighedDistance(q_a, q_b) (steeringMethod::Straight) is computed there, so that a call
   moves 16 B per edge to the device instead of 2 nq + 1 doubles.  Outputs as hppb200_validate_edges, plus length [n]
   (may be NULL).  Host arrays only; `nn` and `ws` must share device and robot. */
int hppb200_validate_edges_indexed(hppb200_workspace* ws, hppb200_knn* nn, int method, double param, int reverse,
                                   const int64_t* node_a, const int64_t* node_b, size_t n, uint32_t flags,
                                   uint8_t* valid, double* t_fail, double* t_last_valid, int32_t* pair,
                                   uint8_t* reason, double* length);

/* Exchange helpers of the one-process-per-GPU form (bench.py --gpus N, hpp-core_b200/sharding.py), where the path's only
   cross-GPU data are validity masks and k-NN candidates.  DEVICE pointers, asynchronous on
   `cuda_stream` (a cudaStream_t; the caller's NCCL collective goes on the same stream).
   pack: valid [n] (0 / 1) -> bits [(n + 31) / 32], bit (i % 32) of word (i / 32), so that the all-gather moves n / 8
   bytes per rank; unpack is the inverse.
   merge: dist / idx [nshards][nqueries][k] = the all-gathered per-shard results of hppb200_knn_query with GLOBAL node
   indices (-1 padded), each list nearest first -> the k nearest over all shards by the key (distance, node index),
   i.e. the reference's tie rule (strict '<', src/nearest-neighbor/basic.cc:66,106) carried across shards. */
int hppb200_pack_mask(void* cuda_stream, const uint8_t* valid, size_t n, uint32_t* bits);
int hppb200_unpack_mask(void* cuda_stream, const uint32_t* bits, size_t n, uint8_t* valid);
int hppb200_knn_merge_shards(void* cuda_stream, const double* dist, const int64_t* idx, int nshards,
                             size_t nqueries, int k, int64_t* out_idx, double* out_dist);

/* ---- fused exchange step over NVLink peer memory (processes of one node) ----
   The sharded form's only exchange is the gather of the validity masks (the per-unit results of ConfigValidations /
   PathValidation::validate over disjoint index ranges).  hppb200_exchange_mask does it in ONE kernel: the
   rank's mask is bit-packed and stored straight into every rank's gather buffer (cudaIpc-mapped peer memory), the last
   CTA publishes an epoch flag on every rank, and a one-warp kernel on the same stream waits for all ranks' flags.  No
   counterpart in the reference (hpp-core is single-process); it replaces hppb200_pack_mask + an NCCL all-gather.
   create: allocates this rank's buffer for masks of up to max_units units and returns its 64-byte IPC handle; the caller
   exchanges the handles of all ranks (any transport) and passes them, rank-major, to open_peers (which blocks until the
   mappings exist: every rank must have called create first).  exchange_mask: collective, same n on every rank, asynchronous
   on cuda_stream; *gathered points to world * ceil(n / 32) words in THIS rank's memory, rank r's words at r * ceil(n / 32),
   valid until the second next call (two slots by call parity). */
typedef struct hppb200_exchange hppb200_exchange;
int hppb200_exchange_create(int device, int rank, int world, size_t max_units, hppb200_exchange** out, unsigned char* handle64);
int hppb200_exchange_open_peers(hppb200_exchange* ex, const unsigned char* handles /* [world][64] */);
int hppb200_exchange_mask(hppb200_exchange* ex, void* cuda_stream, const uint8_t* valid, size_t n, const uint32_t** gathered);
/* Synchronises the device; *timed_out != 0 when a wait of an earlier hppb200_exchange_mask gave up after ~2 s (a peer never
   delivered): the gathered words of that call are not complete. */
int hppb200_exchange_status(hppb200_exchange* ex, int* timed_out);
void hppb200_exchange_destroy(hppb200_exchange* ex);

/* ---- batched planner front-end: host-side array passes (no device work) ---- */
/* kPrmStar::linkNodes / connectNodeToClosestNeighbors (reference src/path-planner/k-prm-star.cc:134-193) over the neighbour
   lists of ALL nodes at once.  nbrs / dist [n][k]: the k nearest nodes of every node, nearest first (-1 padded), as
   hppb200_knn_query returns them with the nodes themselves as queries.  The reference walks node i's list and skips a
   neighbour j that already has an edge j -> i; so a directed candidate (i, j) is
     kind 1: validated whatever happens (j > i, or i is not among j's neighbours);
     kind 2: validated only if the mirrored candidate (j, i) was found invalid; mirror[slot] = flat slot j * k + position
             of i in j's list, where that outcome lives;
     kind 0: no candidate (padding, the node itself).
   symmetric_distance != 0: the distance is symmetric bit for bit (WeighedDistance on a vector space, reference
   src/weighed-distance.cc:272-289) and the lists are sorted by (distance, index): membership by binary search on that key;
   otherwise by scanning j's list. */
int hppb200_kprm_link_candidates(size_t n, int k, const int64_t* nbrs, const double* dist, int symmetric_distance,
                                 uint8_t* kind /* [n][k] */, int64_t* mirror /* [n][k] */);
/* Roadmap::addEdges + ConnectedComponent::merge (reference src/roadmap.cc:209-255) for `nedges` edges (pairs [nedges][2],
   node indices) in insertion order: the component of an edge's second node joins the component of its first node.
   labels [n] in/out: connected-component label per node (nodes sharing a label on entry are one component). */
int hppb200_roadmap_merge_edges(size_t n, int64_t* labels, const int64_t* pairs, size_t nedges);

/* ---- nearest neighbour (WeighedDistance brute force) ------------------- */
int hppb200_knn_create(const hppb200_model* m, int device, hppb200_knn** out);
void hppb200_knn_destroy(hppb200_knn* nn);
/* Append nodes (roadmap insertion order = scan order); cc [n] connected-component
   label per node (may be NULL = 0). */
int hppb200_knn_append(hppb200_knn* nn, const double* q, const int32_t* cc, size_t n);
int hppb200_knn_relabel(hppb200_knn* nn, int32_t from_cc, int32_t to_cc);
/* All connected-component labels at once, cc [n], n == hppb200_knn_size (many ConnectedComponent::merge calls of a
   batched planner step collapsed into their final labelling; reference src/roadmap.cc:209-255). */
int hppb200_knn_set_labels(hppb200_knn* nn, const int32_t* cc, size_t n);
size_t hppb200_knn_size(const hppb200_knn* nn);
/* queries [nqueries][nq]; cc < 0 searches the whole roadmap.  idx [nqueries][k]
   nearest first (-1 padded), dist [nqueries][k]. */
int hppb200_knn_query(hppb200_knn* nn, const double* queries, size_t nqueries,
                      int k, int32_t cc, uint32_t flags, int64_t* idx, double* dist);
/* hppb200_knn_query takes the tensor-core candidate pass (tcgen05 TF32 GEMM of the query tile against the node
   stream + exact FP64 re-rank of the survivors; same indices and distances as the exact scan, bit for bit) when
   the model is a vector space (no free-flyer, 3 nq + 6 <= 48), k <= 256 (a component filter cc >= 0 is applied by
   the exact re-rank), the mirror holds >= 16384 nodes
   and the call carries >= 32 queries (env HPPB200_KNN_TC = 0 disables it, HPPB200_KNN_TC_MIN sets the query
   count); otherwise the exact FP64 scan.  Both replace nearestNeighbor::Basic::KnearestSearch
   (reference src/nearest-neighbor/basic.cc:84-126).
   Statistics of the last query: out[0] 1 if the tensor-core pass ran, out[1] prefix stages, out[2] queries redone
   by the exact scan (candidate list overflow), out[3] candidate groups (8 consecutive nodes each) re-ranked in
   the last stage. */
int hppb200_knn_tc_stats(hppb200_knn* nn, int64_t* out /* [4] */);
/* Device time (CUDA events on the mirror's stream) of the whole-roadmap tcgen05 filter pass of the last query and
   the (query, node) pairs it evaluated (padded to tiles): the kernel the k-NN roofline is reported on. */
int hppb200_knn_tc_filter_time(hppb200_knn* nn, double* ms, double* pairs);
/* Device time (CUDA events on the mirror's stream) of the single-query kernel of the last hppb200_knn_query call
   that took the one-kernel path (one query in host memory, k <= 32; kernels_knn_latency.cu), and the bytes of the node
   stream it scanned; 0 / 0 when the last call took another path.  The events are recorded only when the mirror was
   created with HPPB200_KNN_TIMING=1 in the environment (two event records per call are not free at ~20 us). */
int hppb200_knn_latency_time(hppb200_knn* nn, double* ms, double* bytes);
/* Test hook: approximate squared distances (the TMEM accumulator) of up to 128 queries against the first 256
   nodes, out [128][256] row-major FP32; swap_offsets != 0 exchanges the descriptor byte offsets (bring-up). */
int hppb200_knn_tc_debug_tile(hppb200_knn* nn, const double* queries, size_t nqueries, int swap_offsets,
                              float* out);
/* count [nqueries]; idx [nqueries][cap] in scan order (strict d < radius). */
int hppb200_knn_within_ball(hppb200_knn* nn, const double* queries, size_t nqueries,
                            double radius, int32_t cc, size_t cap, int64_t* idx,
                            int64_t* count);

#ifdef __cplusplus
}
#endif
#endif /* HPPB200_H */
