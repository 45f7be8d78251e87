// Single-query nearest-neighbour path (kernels_knn_latency.cu)
#pragma once
#include "knn_state.hpp"

namespace hppb200 {

constexpr int KNN_LAT_MAX_NQ = 64;  // the query travels in the kernel's parameter block
struct LatQuery {
  double q[KNN_LAT_MAX_NQ];
};

bool knnLatencyEligible(const hppb200_knn* nn, size_t nqueries, int k);
void knnLatencyRun(hppb200_knn* nn, const double* h_query, int k, int32_t cc, long long* outIdx, double* outDist);

}  // namespace hppb200
