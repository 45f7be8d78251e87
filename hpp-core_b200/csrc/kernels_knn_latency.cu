// Single-query nearest neighbour: the planner's one-item call (BiRRTPlanner::oneStep asks for ONE nearest node per
// extension, reference src/bi-rrt-planner.cc:129-160 -> Roadmap::nearestNode -> nearestNeighbor::Basic::search,
// src/nearest-neighbor/basic.cc:52-118).
//
// The whole call is ONE kernel and one stream synchronisation:
//   * the query travels in the kernel's parameter block (no host-to-device copy to wait for),
//   * the grid scans the node stream once (48 B per node at nq = 6: HBM-bound), every warp keeping the k smallest
//     keys (distance, node index) sorted across its lanes, warps merged through shared memory, one list per CTA,
//   * the LAST CTA to finish (ticket counter, __threadfence) merges the per-CTA lists and writes the k results
//     straight into the caller's pinned result block (mapped host memory): no second launch, no device-to-host copy.
// Keys and distance arithmetic are those of the exact scan (kernels_knn.cu, dev_knn.cuh): same bits, same tie rule
// (equal distances: the earlier node, basic.cc:66,106).
#include <algorithm>
#include <cstdlib>
#include <climits>

#include "dev_model.cuh"
#include "dev_knn.cuh"
#include "knn_latency.hpp"

namespace hppb200 {
void setLastError(const std::string& s);

constexpr int LAT_THREADS = 256;
constexpr int LAT_WARPS = LAT_THREADS / 32;

template <bool FF>
__global__ void __launch_bounds__(LAT_THREADS)
knnLatencyKernel(ModelView mv, const double* __restrict__ nodes, const int32_t* __restrict__ cc, size_t n, LatQuery Q, int32_t ccFilter,
                 int k, Key* __restrict__ partial, unsigned* __restrict__ ticket, long long* __restrict__ outIdx, double* __restrict__ outDist) {
  const Tables T = makeTables(mv.blob, mv);  // (global memory: staging the tables would cost more than the few reads)
  __shared__ double sq[KNN_LAT_MAX_NQ];
  __shared__ double sw2[KNN_LAT_MAX_NQ];  // njoints <= nq
  __shared__ Key lists[LAT_WARPS][32];
  __shared__ bool isLast;
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int x = threadIdx.x; x < mv.nq; x += LAT_THREADS) sq[x] = Q.q[x];
  for (int j = threadIdx.x; j < mv.njoints; j += LAT_THREADS) { const double w = T.weights[j]; sw2[j] = w * w; }
  __syncthreads();

  // ---- scan: CTA b takes the contiguous slice [n b / grid, n (b + 1) / grid), a warp 32 consecutive nodes per trip ----
  const size_t lo = (size_t)((unsigned long long)n * blockIdx.x / gridDim.x);
  const size_t hi = (size_t)((unsigned long long)n * (blockIdx.x + 1) / gridDim.x);
  Key mine = {CUDART_INF, LLONG_MAX}, kth = mine;
  for (size_t base = lo + (size_t)warp * 32; base < hi; base += (size_t)LAT_WARPS * 64) {
    // two trips per iteration: both nodes' loads are in flight before the first insertion logic runs
    const size_t i0 = base + lane, i1 = base + (size_t)LAT_WARPS * 32 + lane;
    const bool ok0 = i0 < hi && (ccFilter < 0 || cc[i0] == ccFilter);
    const bool ok1 = i1 < hi && (ccFilter < 0 || cc[i1] == ccFilter);
    double d0[1], d1[1];
    weighedDistanceBatch<1, FF>(T, mv, nodes + (ok0 ? i0 : lo) * (size_t)mv.nq, sq, sw2, d0);
    weighedDistanceBatch<1, FF>(T, mv, nodes + (ok1 ? i1 : lo) * (size_t)mv.nq, sq, sw2, d1);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const Key c = {u ? d1[0] : d0[0], (long long)(u ? i1 : i0)};
      unsigned m = __ballot_sync(FULL, (u ? ok1 : ok0) && keyLess(c, kth));
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const Key x = {__shfl_sync(FULL, c.d, src), __shfl_sync(FULL, c.i, src)};
        if (keyLess(x, kth)) topkInsert(mine, kth, x, k, lane);  // (warp-uniform condition)
      }
    }
  }
  // ---- warps of the CTA -> one list ----
  lists[warp][lane] = mine;
  __syncthreads();
  if (warp == 0) {
    Key acc = lists[0][lane];
    Key ak = {__shfl_sync(FULL, acc.d, k - 1), __shfl_sync(FULL, acc.i, k - 1)};
    for (int w = 1; w < LAT_WARPS; ++w) {
      const Key c = lists[w][lane];
      unsigned m = __ballot_sync(FULL, lane < k && keyLess(c, ak));
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const Key x = {__shfl_sync(FULL, c.d, src), __shfl_sync(FULL, c.i, src)};
        if (keyLess(x, ak)) topkInsert(acc, ak, x, k, lane);
      }
    }
    if (lane < k) partial[(size_t)blockIdx.x * k + lane] = acc;
    __threadfence();  // the list is visible device-wide before the ticket is taken
    __syncwarp();
    if (lane == 0) isLast = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!isLast) return;
  __threadfence();

  // ---- last CTA: the per-CTA lists are one flat stream of grid * k keys; warp w merges its share, warp 0 the 8 results ----
  const int total = (int)gridDim.x * k;
  Key acc = {CUDART_INF, LLONG_MAX}, ak = acc;
  for (int f0 = warp * 64; f0 < total; f0 += LAT_WARPS * 64) {
    Key c[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int f = f0 + u * 32 + lane;
      c[u].d = CUDART_INF; c[u].i = LLONG_MAX;
      if (f < total) {  // (written by other CTAs of this launch: read past L1)
        const double2 v = __ldcg(reinterpret_cast<const double2*>(partial) + f);
        c[u].d = v.x;
        c[u].i = __double_as_longlong(v.y);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      unsigned m = __ballot_sync(FULL, keyLess(c[u], ak));
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const Key x = {__shfl_sync(FULL, c[u].d, src), __shfl_sync(FULL, c[u].i, src)};
        if (keyLess(x, ak)) topkInsert(acc, ak, x, k, lane);
      }
    }
  }
  __syncthreads();  // (lists[] is reused)
  lists[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) {
    acc = lists[0][lane];
    ak.d = __shfl_sync(FULL, acc.d, k - 1);
    ak.i = __shfl_sync(FULL, acc.i, k - 1);
    for (int w = 1; w < LAT_WARPS; ++w) {
      const Key c = lists[w][lane];
      unsigned m = __ballot_sync(FULL, lane < k && keyLess(c, ak));
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const Key x = {__shfl_sync(FULL, c.d, src), __shfl_sync(FULL, c.i, src)};
        if (keyLess(x, ak)) topkInsert(acc, ak, x, k, lane);
      }
    }
    if (lane < k) {
      outIdx[lane] = acc.i != LLONG_MAX ? acc.i : -1;  // fewer than k candidates: -1 / +inf, like the exact scan
      outDist[lane] = acc.d;
    }
    if (lane == 0) *ticket = 0;  // ready for the next call on this stream
  }
}


bool knnLatencyEligible(const hppb200_knn* nn, size_t nqueries, int k) {
  return nqueries == 1 && k >= 1 && k <= 32 && nn->model->nq <= KNN_LAT_MAX_NQ && nn->n > 0 && nn->lat_enabled;
}

// one query (host memory) -> idx / dist [k] (pinned, device-accessible host memory or device memory); nn->stream, not synchronised
void knnLatencyRun(hppb200_knn* nn, const double* h_query, int k, int32_t cc, long long* outIdx, double* outDist) {
  LatQuery Q;
  const int nq = nn->model->nq;
  for (int x = 0; x < nq; ++x) Q.q[x] = h_query[x];
  // grid: a few CTAs per SM (enough loads in flight for the HBM stream), never more than one warp trip pair per CTA wants
  const size_t perCta = (size_t)LAT_WARPS * 64;
  static const size_t ctasPerSm = [] { const char* v = std::getenv("HPPB200_KNN_LAT_CTAS"); return v ? (size_t)std::max(1, std::atoi(v)) : (size_t)2; }();  // measured per SM: 1: 35 us, 2: 28.5, 4: 30.5, 8: 38, 16: 49 us (k = 1; the last CTA merges grid x k keys)
  const unsigned grid = (unsigned)std::max<size_t>(1, std::min<size_t>((size_t)nn->num_sms * ctasPerSm, (nn->n + perCta - 1) / perCta));
  if (!nn->lat_ticket_zeroed) {  // sized for the largest grid once: the ticket counter must survive from call to call
    nn->d_lat.reserve((size_t)nn->num_sms * ctasPerSm * 32 * sizeof(Key) + 256);
    HPPB200_CUDA_CHECK(cudaMemsetAsync(nn->d_lat.p, 0, 256, nn->stream));
    nn->lat_ticket_zeroed = true;
  }
  unsigned* ticket = (unsigned*)nn->d_lat.p;
  Key* partial = (Key*)(nn->d_lat.p + 256);
  auto kern = nn->ff ? knnLatencyKernel<true> : knnLatencyKernel<false>;
  if (nn->lat_timing) {
    if (!nn->lat_ev0) { HPPB200_CUDA_CHECK(cudaEventCreate(&nn->lat_ev0)); HPPB200_CUDA_CHECK(cudaEventCreate(&nn->lat_ev1)); }
    HPPB200_CUDA_CHECK(cudaEventRecord(nn->lat_ev0, nn->stream));
  }
  kern<<<grid, LAT_THREADS, 0, nn->stream>>>(nn->view, nn->d_nodes, nn->d_cc, nn->n, Q, cc, k, partial, ticket, outIdx, outDist);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  if (nn->lat_timing) HPPB200_CUDA_CHECK(cudaEventRecord(nn->lat_ev1, nn->stream));
}

}  // namespace hppb200

extern "C" int hppb200_knn_latency_time(hppb200_knn* nn, double* ms, double* bytes) {
  if (!nn || !ms || !bytes) { hppb200::setLastError("null argument"); return HPPB200_ERR_INVALID_ARGUMENT; }
  *ms = 0.0;
  *bytes = 0.0;
  if (!nn->lat_last_used || !nn->lat_timing || !nn->lat_ev0) return HPPB200_OK;
  float f = 0;
  if (cudaSetDevice(nn->device) != cudaSuccess || cudaEventElapsedTime(&f, nn->lat_ev0, nn->lat_ev1) != cudaSuccess) {
    hppb200::setLastError("CUDA: event timing of the single-query kernel failed");
    return HPPB200_ERR_CUDA;
  }
  *ms = (double)f;
  *bytes = (double)nn->n * (double)nn->model->nq * sizeof(double);
  return HPPB200_OK;
}
