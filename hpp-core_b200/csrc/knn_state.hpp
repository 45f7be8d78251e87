// State of the device roadmap mirror (hppb200_knn) shared by the exact FP64 scan (kernels_knn.cu) and the
// tensor-core candidate pass (kernels_knn_tc.cu).
#pragma once
#include "engine.hpp"

struct hppb200_knn {
  const hppb200::Model* model = nullptr;
  int device = 0;
  hppb200::ModelView view;
  cudaStream_t stream = nullptr;
  double* d_nodes = nullptr;  // [cap][nq]
  int32_t* d_cc = nullptr;
  size_t n = 0, cap = 0;
  hppb200::DeviceBuffer<double> d_queries, d_dist;
  hppb200::DeviceBuffer<long long> d_idx, d_count;
  hppb200::DeviceBuffer<unsigned char> d_keys;  // previous keys [Q] + per-tile partial keys [tiles][Q]
  int num_sms = 148;
  unsigned char* h_stage = nullptr;  // pinned staging block for small query calls (mapped: d_stage_map is its device address)
  unsigned char* d_stage_map = nullptr;
  // single-query path (kernels_knn_latency.cu): ticket counter (first 256 B) + per-CTA candidate lists
  hppb200::DeviceBuffer<unsigned char> d_lat;
  bool lat_ticket_zeroed = false;
  bool lat_timing = false;  // HPPB200_KNN_TIMING=1: events around the single-query kernel (hppb200_knn_latency_time)
  bool lat_last_used = false;
  cudaEvent_t lat_ev0 = nullptr, lat_ev1 = nullptr;
  bool lat_enabled = true;  // HPPB200_KNN_LATENCY (0: single queries go through the tiled scan as before)

  // ---- tensor-core candidate pass (kernels_knn_tc.cu) ----
  bool ff = false;           // the model has free-flyer joints (no GEMM form of the distance)
  int tc_kc = 0;             // K = 8 * tc_kc operand slots per node (0: model not eligible)
  int tc_min_queries = 32;   // HPPB200_KNN_TC_MIN
  bool tc_prefilter = true;  // HPPB200_KNN_TC_PREFILTER (0 switches it off): k <= 32 re-rank drops keys above the threshold before insertion
  double tc_ratio = 16.0;    // HPPB200_KNN_TC_RATIO: prefix growth per stage
  bool tc_enabled = true;    // HPPB200_KNN_TC
  float* d_tc_nodes = nullptr;   // node operand tiles in the UMMA K-major canonical layout, [tile][256 x 8 kc] tf32
  size_t tc_cap_tiles = 0;       // tiles allocated
  size_t tc_n = 0;               // nodes already converted
  hppb200::DeviceBuffer<float> d_tc_queries, d_tc_thr;
  hppb200::DeviceBuffer<double> d_tc_scale, d_tc_kdist;
  hppb200::DeviceBuffer<int> d_tc_cand, d_tc_cnt;
  hppb200::DeviceBuffer<unsigned> d_tc_misc;  // [0] max node norm (float bits), [1] overflowed queries of the last stage
  hppb200::DeviceBuffer<long long> d_tc_kidx;
  hppb200::DeviceBuffer<int> d_tc_flagged;
  // statistics of the last tensor-core query (hppb200_knn_tc_stats)
  unsigned long long tc_last_candidates = 0, tc_last_overflow = 0;
  int tc_last_stages = 0;
  size_t tc_last_rows = 0;
  int tc_last_gused = 0;
  bool tc_last_used = false;
  double tc_last_filter_ms = 0, tc_last_filter_pairs = 0;  // whole-roadmap filter pass of the last query (CUDA events)
  cudaEvent_t tc_ev0 = nullptr, tc_ev1 = nullptr;
};

namespace hppb200 {

struct KnnKey {
  double d;
  long long i;
};

// exact FP64 scan over the first `nNodes` nodes of the mirror (kernels_knn.cu); everything on nn->stream
void knnExactRun(hppb200_knn* nn, const double* d_q, size_t nqueries, int k, int32_t cc, size_t nNodes, long long* d_i,
                 double* d_d);
// tensor-core candidate pass + exact FP64 re-rank (kernels_knn_tc.cu); same results as knnExactRun over all nodes
bool knnTcEligible(const hppb200_knn* nn, size_t nqueries, int k, int32_t cc);
void knnTcRun(hppb200_knn* nn, const double* d_q, size_t nqueries, int k, int32_t cc, long long* d_i, double* d_d);
void knnTcInit(hppb200_knn* nn);
void knnTcDestroy(hppb200_knn* nn);

}  // namespace hppb200
