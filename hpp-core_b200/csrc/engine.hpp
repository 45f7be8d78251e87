// Host side of the engine: the model compiler (flattened robot/scene tables in the
// reference's pair ORDER), the device blob layout shared with the kernels, and the
// workspace (one CUDA stream + scratch).  See DESIGN.md §2-§3.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/hppb200.h"

namespace hppb200 {

// ---- records inside the device blob (all 8-byte aligned, blob 16-byte padded) ----------
struct DevShapeH {  // mirrors DevShape in dev_narrow.cuh
  double p[4];
  int32_t type, vbegin, vcount, pad;
};
struct DevJoint {
  int32_t parent, type, idx_q, idx_v;
  double P[12];
};
// Bounding core used by the conservative cull test: a point (or, for a
// capsule, its axis segment) at local offset bc, inflated by br.  br is rounded UP.
struct DevBound {
  double bc[3];
  double br;
};
struct DevGeom {
  int32_t joint, pad;
  double P[12];
  DevShapeH shape;
  DevBound bound;
};
struct DevObstacle {
  double P[12];
  DevShapeH shape;
  DevBound bound;
  double wc[3];  // world position of the bounding core (R bc + t), fixed
  double pad;
};
// a/b: >= 0 robot geometry index, < 0 obstacle -(k+1); already canonically ordered
// (shape type of a <= shape type of b) so that the kernel can dispatch on `kind` directly.
struct DevPair {
  int32_t a, b, kind;
  int32_t bound;   // 0: point-point, 1: segment(a)-point(b) cull test
  double margin;
  double cull2;    // squared core distance above which the pair provably does not collide
};
struct DevBodyPair {
  int32_t joint_a, joint_b;
  int32_t pair_begin, pair_end;    // into bodyGeomPairs
  int32_t coef_begin, coef_end;    // a -> b
  int32_t rcoef_begin, rcoef_end;  // b -> a (empty when joint_b == 0)
  int32_t has_gjk;                 // some geometry pair of this body pair goes through GJK
};
struct DevCoef {
  int32_t idx_v;
  int16_t nv, idx_q;  // velocity segment of the joint and where its configuration entries start
  double value;
};

// Offsets (bytes) of each table in the blob + counts; passed to kernels by value.
struct ModelView {
  const unsigned char* blob;  // global memory
  const double* vertices;     // global memory convex vertex pool (not staged)
  uint32_t blob_bytes;
  uint32_t edge_blob_bytes;  // prefix without the configuration pair table (continuous validation)
  uint32_t off_joints, off_geoms, off_obstacles, off_cpairs, off_bodypairs, off_bgpairs, off_coefs, off_rcoefs,
      off_lower, off_upper, off_weights;
  int32_t njoints, ngeoms, nobstacles, ncpairs, nbodypairs, nbgpairs, ncoefs, nrcoefs, nq, nv;
};

struct BodyPairHost {
  int joint_a, joint_b;
  std::vector<DevPair> pairs;      // reference order (first = robot geometry)
  std::vector<std::pair<int, double>> coefs, rcoefs;  // (child joint, value)
};

struct Model {
  explicit Model(const hppb200_model_desc& d);
  ~Model();
  int njoints = 0, nq = 0, nv = 0;
  std::vector<DevJoint> joints;       // index 0 = joint 1
  std::vector<int> joint_nq, joint_nv;
  std::vector<DevGeom> geoms;
  std::vector<DevObstacle> obstacles;
  std::vector<double> lower, upper, weights, margins, vertices, radii;
  double break_distance = 1e-2;
  // reference-order tables (for the query API and reports)
  std::vector<DevPair> cPairs;        // first/second as in cPairs_ (NOT canonicalised); pPairs_ appended when checked
  std::vector<DevPair> pPairs, dPairs;  // parameterized / disabled pairs (ObstacleUser::filterCollisionPairs)
  int nParameterized = 0;             // trailing entries of cPairs that are pPairs_
  std::vector<DevPair> dboPairs;      // DistanceBetweenObjects pairs (inner object, obstacle), canonical order, margin 0
  std::vector<BodyPairHost> bodyPairs;
  // device
  std::vector<unsigned char> blob;
  ModelView view{};                   // blob/vertices pointers are filled per device
  ModelView deviceView(int device) const;  // uploads on first use

 private:
  void buildConfigPairs(const hppb200_model_desc& d);
  void buildBodyPairs(const hppb200_model_desc& d);
  void applyPairOps(const hppb200_model_desc& d);
  void finishBodyPair(BodyPairHost& bp) const;
  void buildBlob();
  double margin(int j1, int j2) const { return margins.empty() ? 0.0 : margins[(size_t)j1 * (njoints + 1) + j2]; }
  mutable std::mutex mu_;
  mutable std::map<int, std::pair<void*, void*>> dev_;  // device -> (blob, vertices)
};

struct DeviceStats {  // accumulated by kernels with atomics
  unsigned long long configs, pair_tests, gjk_iterations, edge_samples, iter_cap_hits, gjk_cap_hits, overflow;
  unsigned long long engine_ticks, discarded_tests;  // continuous engine: CTA ticks, speculative tests thrown away
};

template <typename T>
struct DeviceBuffer {
  T* p = nullptr;
  size_t cap = 0;
  void reserve(size_t n) {
    if (n <= cap) return;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = n + n / 8 + 64;
    if (cudaMalloc((void**)&p, want * sizeof(T)) != cudaSuccess) throw std::runtime_error("cudaMalloc failed");
    cap = want;
  }
  ~DeviceBuffer() { if (p) cudaFree(p); }
};

constexpr size_t SMALL_STAGE_BYTES = 64 * 1024;
struct Workspace {
  Workspace(const Model* m, int device);
  ~Workspace();
  const Model* model;
  int device;
  int num_sms = 148;
  cudaStream_t stream = nullptr;
  bool own_stream = true;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaStream_t stream2 = nullptr, stream3 = nullptr, stream4 = nullptr, stream5 = nullptr;  // H2D / D2H streams of the sliced host-pointer path (created on first use)
  std::vector<cudaEvent_t> evSlice;
  cudaEvent_t evFork = nullptr, evJoin = nullptr;
  ModelView view;
  DeviceStats* d_stats = nullptr;
  hppb200_stats stats{};
  uint32_t max_iterations = 1u << 20;
  bool use_engine = true;      // Progressive: control/narrow-phase engine (HPPB200_ENGINE=0 forces the walker)
  int engine_threads = 128;    // CTA size of the engine (HPPB200_ENGINE_THREADS: 64 | 96 | 128)
  int engine_kind = 2;         // Progressive: 2 wavefront rounds (kernels_wavefront.cu), 1 CTA engine (HPPB200_ENGINE)
  size_t wavefront_min = 100000;  // smallest batch that takes the wavefront engine (HPPB200_WF_MIN)
  long wavefront_tail = -1;    // slots left when the wavefront hands the tail to the walker (-1: min(2048, n/64); HPPB200_WF_TAIL)
  size_t wavefront_slots = 0;  // resident edge slots of the wavefront engine (0: 2048 per SM; HPPB200_WF_SLOTS)
  double engine_alpha = 1.2;   // candidate prediction: alpha x previous step (HPPB200_ENGINE_ALPHA; 1.2 measured best)
  // staging for host-pointer calls
  DeviceBuffer<double> d_q, d_q2, d_L, d_tfail, d_tlast, d_scratch;
  DeviceBuffer<uint8_t> d_valid, d_reason, d_reason_int;
  DeviceBuffer<uint32_t> d_list, d_cont;
  DeviceBuffer<double> d_sv_dbl;
  DeviceBuffer<uint16_t> d_sv_order;
  DeviceBuffer<uint8_t> d_sv_flags;
  DeviceBuffer<unsigned long long> d_count;
  DeviceBuffer<int32_t> d_pair, d_pair_int, d_chain, d_cpair, d_cbad;
  DeviceBuffer<uint8_t> d_cvalid, d_creason;
  DeviceBuffer<long long> d_ia, d_ib;
  DeviceBuffer<DevPair> d_dbo;
  bool dboUploaded = false;
  DeviceBuffer<uint32_t> d_mask;
  DeviceBuffer<int32_t> d_ncoll;
  DeviceBuffer<double> d_ctfail, d_ctlast;
  DeviceBuffer<unsigned long long> d_jobs;
  unsigned long long last_jobs = 0;
  // first-sample screen of the edge path (kernels_edges.cu::launchValidateEdges)
  bool first_sample_screen = true;     // HPPB200_FIRST_SAMPLE_SCREEN=0 switches it off
  size_t first_sample_min = 2048;      // smallest batch that is screened (HPPB200_FS_MIN)
  bool first_sample_order = true;      // the screened edges go to the engines closest-to-an-obstacle first (HPPB200_FS_ORDER=0: input order)
  DeviceBuffer<double> d_fs_q, d_fs_q1, d_fs_q2, d_fs_L, d_fs_tfail, d_fs_tlast;
  DeviceBuffer<uint8_t> d_fs_valid, d_fs_reason, d_fs_ovalid, d_fs_oreason;
  DeviceBuffer<int32_t> d_fs_first, d_fs_opair;
  DeviceBuffer<uint32_t> d_fs_list;
  DeviceBuffer<unsigned long long> d_fs_cnt, d_fs_dmin, d_fs_keys, d_fs_keys2;
  DeviceBuffer<uint32_t> d_fs_list2;
  DeviceBuffer<unsigned char> d_fs_tmp;
  unsigned long long* h_fs = nullptr;  // pinned: survivor count + the chunks' queue counters
  // host-pointer edge calls (capi.cu): q2 and L arrive on their own stream while the screen runs on q1; launchValidateEdges waits for
  // this event (recorded behind those copies) before it first touches q2 / L, and clears it
  cudaStream_t edge_copy_stream = nullptr;
  cudaEvent_t ev_edge_in = nullptr, ev_edge_q1 = nullptr;
  cudaEvent_t edge_inputs_ready = nullptr;  // non-owning: ev_edge_in for the duration of a call, else null
  // one-item / small host-pointer calls (a planner's single-item virtuals): inputs packed in ONE pinned block and one H2D, outputs
  // in one device block and one D2H (capi.cu)
  unsigned char* h_small = nullptr;       // pinned, SMALL_STAGE_BYTES
  DeviceBuffer<unsigned char> d_small;
  bool overflow_known = false;            // DeviceStats::overflow as of the end of the last edge call, when that call read it last
  unsigned long long overflow_cached = 0;
  unsigned long long* h_njobs = nullptr;  // pinned: queue counter of the last configuration batch
  DeviceBuffer<unsigned char> d_state, d_tail, d_sort;
  bool wavefront_lpt = true;   // wavefront engines pull the edges longest first (kernels_sort.cu; HPPB200_WF_LPT=0: input order)
  bool use_tail = true;        // Dichotomy: CTA-per-edge pass for handed-over edges (HPPB200_TAIL=0: walker large-list pass only)
  bool tail_smem = true;       // first CTA-per-edge pass keeps the per-pair lists in shared memory (HPPB200_TAIL_SMEM)
  size_t tail_max = ~size_t(0);      // Dichotomy batches up to this size that the wavefront engine does not take go to the CTA-per-edge kernel (HPPB200_TAIL_MAX)
  size_t prog_tail_max = 4096;         // Progressive batches up to this size go to the CTA-per-edge kernel (HPPB200_PROG_TAIL_MAX)
  void syncStats();
};

void setLastError(const std::string& s);
double measureFp64Peak(int device);

// kernel launchers (kernels_config.cu / kernels_edges.cu), all asynchronous on ws->stream
void finishValidateConfigs(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid, int32_t* d_first,
                           uint8_t* d_reason);
void launchValidateConfigs(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid,
                           int32_t* d_first, uint8_t* d_reason);
void noallocSetup(int device);  // resolves the static scratch array's device address (cudaGetSymbolAddress), once per device
void launchValidateConfigsNoAlloc(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid, int32_t* d_first,
                                  uint8_t* d_reason);
size_t firstSampleChunk();
size_t firstSampleQueueCap(size_t n);
bool firstSampleScreenAvailable(Workspace* ws);
bool launchFirstSampleScreen(Workspace* ws, const double* d_q, size_t n, uint8_t* d_valid, int32_t* d_first, uint8_t* d_reason,
                             unsigned long long* d_njobs, unsigned long long* d_dmin);
void launchValidateConfigsSmall(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid, int32_t* d_first, uint8_t* d_reason);
void launchValidateConfigsHost(Workspace* ws, const double* h_q, size_t n, uint32_t flags, uint8_t* h_valid, int32_t* h_first,
                               uint8_t* h_reason);
void launchAllCollisions(Workspace* ws, const double* d_q, size_t n, uint8_t* d_valid, int32_t* d_first, uint32_t* d_mask, int32_t* d_count);
void launchDistanceBetweenObjects(Workspace* ws, const double* d_q, size_t n, double* d_out);
void launchGeometryPlacements(Workspace* ws, const double* d_q, size_t n, double* d_oMg);
void launchPairDistances(Workspace* ws, const double* d_q, size_t n, double* d_dist);
void launchStraightLengths(Workspace* ws, const double* d_q1, const double* d_q2, size_t n, double* d_L);
void launchPathEval(Workspace* ws, const double* d_q1, const double* d_q2, const double* d_L, const double* d_t, size_t n, double* d_q);
bool launchProgressiveEngine(Workspace* ws, double tol, int reverse, const double* d_q1, const double* d_q2,
                             const double* d_L, size_t n, uint8_t* d_valid, double* d_tfail, double* d_tlast,
                             int32_t* d_pair, uint8_t* d_reason, int overflow_reason);
const uint32_t* sortEdgesByLength(Workspace* ws, const double* d_L, size_t n);
bool launchWavefront(Workspace* ws, int method, double tol, int reverse, const double* d_q1, const double* d_q2,
                                const double* d_L, size_t n, uint8_t* d_valid, double* d_tfail, double* d_tlast,
                                int32_t* d_pair, uint8_t* d_reason, int overflow_reason);
// State of wavefront slots (kernels_wavefront.cu) from which the CTA-per-edge kernel continues Progressive edges.
struct TailResume {
  const uint32_t* slots = nullptr;  // the slots to take over (one edge each, at a sample boundary)
  uint32_t E = 0, ordStride = 0;
  const double2* lohi = nullptr;             // [NP][E] validated interval by order position
  const uint16_t* ord = nullptr;             // [E][ordStride] pair id by order position
  const unsigned long long* st_vmask = nullptr;  // [W][E] valid_ bits by order position
  const double* st_t = nullptr;
  const double* st_last = nullptr;
  const uint32_t* st_iters = nullptr;
  const uint32_t* st_edge = nullptr;
  const uint8_t* st_flags = nullptr;         // bits 2-3: `finished`
};
bool launchEdgeTail(Workspace* ws, int method, double tol, int reverse, const double* d_q1, const double* d_q2, const double* d_L,
                    const uint32_t* d_list, size_t n, uint8_t* d_valid, double* d_tfail, double* d_tlast, int32_t* d_pair,
                    uint8_t* d_reason, int overflow_reason, bool smemLists, const TailResume* resume);
bool launchDichotomyTail(Workspace* ws, double tol, int reverse, const double* d_q1, const double* d_q2, const double* d_L,
                         const uint32_t* d_list, size_t n, uint8_t* d_valid, double* d_tfail, double* d_tlast, int32_t* d_pair,
                         uint8_t* d_reason, int overflow_reason, bool smemLists);
void launchChainReduce(Workspace* ws, const int32_t* d_chainBegin, size_t nChains, int reverse, const double* d_L, const uint8_t* eValid,
                       const double* eTfail, const double* eTlast, const int32_t* ePair, const uint8_t* eReason, uint8_t* valid,
                       double* tFail, double* tLast, int32_t* pair, uint8_t* reason, int32_t* firstBad);
void launchValidateSplineEdges(Workspace* ws, int method, double param, int reverse, const double* d_base, const double* d_params,
                               const double* d_L, size_t n, uint32_t flags, uint8_t* d_valid, double* d_tfail, double* d_tlast,
                               int32_t* d_pair, uint8_t* d_reason);
void launchGatherEdgeEnds(cudaStream_t stream, const double* nodes, size_t nNodes, int nq, const int64_t* a, const int64_t* b, size_t n,
                          double* q1, double* q2, unsigned* bad);
void launchPackMask(cudaStream_t stream, const uint8_t* valid, size_t n, uint32_t* bits);
void launchUnpackMask(cudaStream_t stream, const uint32_t* bits, size_t n, uint8_t* valid);
void launchKnnMergeShards(cudaStream_t stream, const double* dist, const int64_t* idx, int W, size_t Q, int k, int64_t* outIdx, double* outDist);
void launchValidateEdges(Workspace* ws, int method, double param, int reverse, const double* d_q1,
                         const double* d_q2, const double* d_L, size_t n, uint32_t flags, uint8_t* d_valid,
                         double* d_tfail, double* d_tlast, int32_t* d_pair, uint8_t* d_reason);

#define HPPB200_CUDA_CHECK(expr)                                                                      \
  do {                                                                                                \
    cudaError_t _e = (expr);                                                                          \
    if (_e != cudaSuccess)                                                                            \
      throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " #expr); \
  } while (0)

}  // namespace hppb200

// opaque C handles (include/hppb200.h)
struct hppb200_model {
  hppb200::Model impl;
  explicit hppb200_model(const hppb200_model_desc& d) : impl(d) {}
};
struct hppb200_workspace {
  hppb200::Workspace impl;
  hppb200_workspace(const hppb200::Model* m, int dev) : impl(m, dev) {}
};
