// Brute-force WeighedDistance nearest neighbour over roadmap nodes mirrored on the device
// (reference src/nearest-neighbor/basic.cc:52-181, src/weighed-distance.cc:272-289).
//
// Exact FP64 path: the node stream (8*nq bytes per node, coalesced) is scanned by the whole GPU in
// tiles and the k smallest are selected by the lexicographic key (distance, node index), which is
// exactly what the reference's linear scan with strict '<' keeps: on equal distances the earlier node
// wins (basic.cc:66,106).  1 < k <= 32: one pass with warp-register top-k lists; k = 1 or k > 32:
// selection passes, each keeping the best key strictly greater than the previous one.
#include <algorithm>
#include <climits>
#include <cstdlib>
#include <cstring>

#include "dev_model.cuh"
#include "dev_knn.cuh"
#include "knn_latency.hpp"

namespace hppb200 {

constexpr int KNN_THREADS = 256;
constexpr size_t KNN_STAGE_BYTES = 256 * 1024;

// ---- tiled scan: the whole GPU works on every query -------------------------------------------------
// Grid (node tiles, query groups): a CTA scans ITS contiguous slice of the node stream once per selection
// pass and evaluates KNN_QB queries against every node it loads (the node stream is read from HBM once per
// query group and pass; the array stays L2 resident across groups and passes when it fits).  Per pass r every
// (tile, query) yields the smallest key strictly greater than the query's previous key; knnReduceKernel
// merges the tiles, appends result r and publishes the new previous key.  Keys are (distance, node index),
// i.e. the reference's strict '<' scan (basic.cc:66,106).  A single query over 1 M nodes is HBM-bound
// (48 B per node at nq = 6); one CTA per query would leave all but one SM idle there.
#ifndef HPPB200_KNN_QB
#define HPPB200_KNN_QB 4
#endif
constexpr int KNN_QB = HPPB200_KNN_QB;

// FF: the model has free-flyer joints (the SO(3) logarithm stays out of the vector-space instantiation)
template <bool FF>
__global__ void __launch_bounds__(KNN_THREADS)
knnTileKernel(ModelView mv, const double* __restrict__ nodes, const int32_t* __restrict__ cc, size_t n,
              const double* __restrict__ queries, int nqueries, int32_t ccFilter, const Key* __restrict__ prev,
              Key* __restrict__ partial) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  unsigned char* sblob = smem + 16;
  stageBlobTMA(sblob, mv.blob, mv.blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);
  double* sq = (double*)(sblob + mv.blob_bytes);  // the CTA's queries and the per-joint w * w
  double* sw2 = sq + KNN_QB * mv.nq;
  for (int x = threadIdx.x; x < KNN_QB * mv.nq; x += KNN_THREADS) {
    const int qq = blockIdx.y * KNN_QB + x / mv.nq;
    sq[x] = queries[(size_t)(qq < nqueries ? qq : nqueries - 1) * mv.nq + x % mv.nq];  // padding repeats the last query
  }
  for (int j = threadIdx.x; j < mv.njoints; j += KNN_THREADS) { const double w = T.weights[j]; sw2[j] = w * w; }
  __syncthreads();
  __shared__ Key red[KNN_QB][KNN_THREADS / 32];
  const int q0 = blockIdx.y * KNN_QB;
  const size_t lo = (size_t)((unsigned long long)n * blockIdx.x / gridDim.x);
  const size_t hi = (size_t)((unsigned long long)n * (blockIdx.x + 1) / gridDim.x);
  Key best[KNN_QB], pv[KNN_QB];
#pragma unroll
  for (int j = 0; j < KNN_QB; ++j) {
    const int q = (q0 + j < nqueries) ? q0 + j : nqueries - 1;  // padding lanes repeat the last query
    pv[j] = prev ? prev[q] : Key{-1.0, -1};  // (first pass: everything is admissible)
    best[j].d = CUDART_INF; best[j].i = LLONG_MAX;
  }
  for (size_t i = lo + threadIdx.x; i < hi; i += KNN_THREADS) {
    if (ccFilter >= 0 && cc[i] != ccFilter) continue;
    const double* node = nodes + i * (size_t)mv.nq;
    double dd[KNN_QB];
    weighedDistanceBatch<KNN_QB, FF>(T, mv, node, sq, sw2, dd);
#pragma unroll
    for (int j = 0; j < KNN_QB; ++j) {
      const Key c = {dd[j], (long long)i};
      if (keyLess(pv[j], c) && keyLess(c, best[j])) best[j] = c;
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < KNN_QB; ++j) {
    Key b = best[j];
    for (int o = 16; o > 0; o >>= 1) {
      const Key other = {__shfl_xor_sync(0xffffffffu, b.d, o), __shfl_xor_sync(0xffffffffu, b.i, o)};
      if (keyLess(other, b)) b = other;
    }
    if (lane == 0) red[j][warp] = b;
  }
  __syncthreads();
  if (threadIdx.x < KNN_QB && q0 + (int)threadIdx.x < nqueries) {
    Key b = red[threadIdx.x][0];
    for (int w = 1; w < KNN_THREADS / 32; ++w)
      if (keyLess(red[threadIdx.x][w], b)) b = red[threadIdx.x][w];
    partial[(size_t)blockIdx.x * nqueries + q0 + threadIdx.x] = b;
  }
}

// one warp per query: merge the tiles of pass r
__global__ void knnReduceKernel(const Key* __restrict__ partial, int ntiles, int nqueries, int k, int r, Key* __restrict__ prev,
                                long long* __restrict__ idx, double* __restrict__ dist) {
  const int q = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (q >= nqueries) return;
  Key b = {CUDART_INF, LLONG_MAX};
  for (int t = lane; t < ntiles; t += 32) {
    const Key c = partial[(size_t)t * nqueries + q];
    if (keyLess(c, b)) b = c;
  }
  for (int o = 16; o > 0; o >>= 1) {
    const Key other = {__shfl_xor_sync(0xffffffffu, b.d, o), __shfl_xor_sync(0xffffffffu, b.i, o)};
    if (keyLess(other, b)) b = other;
  }
  if (lane == 0) {
    const bool found = b.i != LLONG_MAX;
    idx[(size_t)q * k + r] = found ? b.i : -1;
    dist[(size_t)q * k + r] = b.d;  // +inf when fewer than k candidates exist
    prev[q] = b;                    // (inf, LLONG_MAX) admits nothing in later passes
  }
}

__global__ void knnInitPrevKernel(Key* prev, int nqueries) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q < nqueries) { prev[q].d = -1.0; prev[q].i = -1; }
}

// ---- single-pass k selection (k <= 32) ------------------------------------------------------------------
// Every warp keeps, per query, its k smallest keys sorted across the lanes (lane j = j-th smallest, in
// registers).  A node is looked at again only if its key beats the current k-th key (rare after the first few
// hundred nodes); it is then inserted with one ballot + one shuffle.  Warps of a CTA are merged through shared
// memory, CTAs (tiles) by knnMergeTopkKernel.  Keys are (distance, node index), so the result is the reference's
// K-nearest SET (strict '>' against the heap top keeps the earlier node, basic.cc:106) in the engine's
// documented order (ties by node index).
template <bool FF>
__global__ void __launch_bounds__(KNN_THREADS)
knnTileTopkKernel(ModelView mv, const double* __restrict__ nodes, const int32_t* __restrict__ cc, size_t n,
                  const double* __restrict__ queries, int nqueries, int32_t ccFilter, int k, Key* __restrict__ partial) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  unsigned char* sblob = smem + 16;
  stageBlobTMA(sblob, mv.blob, mv.blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);
  double* sq = (double*)(sblob + mv.blob_bytes);  // the CTA's queries and the per-joint w * w
  double* sw2 = sq + KNN_QB * mv.nq;
  for (int x = threadIdx.x; x < KNN_QB * mv.nq; x += KNN_THREADS) {
    const int qq = blockIdx.y * KNN_QB + x / mv.nq;
    sq[x] = queries[(size_t)(qq < nqueries ? qq : nqueries - 1) * mv.nq + x % mv.nq];  // padding repeats the last query
  }
  for (int j = threadIdx.x; j < mv.njoints; j += KNN_THREADS) { const double w = T.weights[j]; sw2[j] = w * w; }
  __syncthreads();
  constexpr int WARPS = KNN_THREADS / 32;
  __shared__ Key lists[WARPS][KNN_QB][32];
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q0 = blockIdx.y * KNN_QB;
  const size_t lo = (size_t)((unsigned long long)n * blockIdx.x / gridDim.x);
  const size_t hi = (size_t)((unsigned long long)n * (blockIdx.x + 1) / gridDim.x);
  Key mine[KNN_QB], kth[KNN_QB];
#pragma unroll
  for (int j = 0; j < KNN_QB; ++j) {
    const int q = (q0 + j < nqueries) ? q0 + j : nqueries - 1;  // padding repeats the last query
    mine[j].d = CUDART_INF; mine[j].i = LLONG_MAX;
    kth[j] = mine[j];
  }
  for (size_t base = lo + (size_t)warp * 32; base < hi; base += (size_t)WARPS * 32) {
    const size_t i = base + lane;
    const bool ok = i < hi && (ccFilter < 0 || cc[i] == ccFilter);
    const double* node = nodes + (ok ? i : lo) * (size_t)mv.nq;
    double dd[KNN_QB];
    weighedDistanceBatch<KNN_QB, FF>(T, mv, node, sq, sw2, dd);
#pragma unroll
    for (int j = 0; j < KNN_QB; ++j) {
      const Key c = {dd[j], (long long)i};
      unsigned m = __ballot_sync(FULL, ok && keyLess(c, kth[j]));
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const Key x = {__shfl_sync(FULL, c.d, src), __shfl_sync(FULL, c.i, src)};
        if (keyLess(x, kth[j])) topkInsert(mine[j], kth[j], x, k, lane);  // (warp-uniform condition)
      }
    }
  }
#pragma unroll
  for (int j = 0; j < KNN_QB; ++j) lists[warp][j][lane] = mine[j];
  __syncthreads();
  if (warp < KNN_QB && q0 + warp < nqueries) {  // warp j merges query j's lists of the other warps into its own
    Key acc = lists[0][warp][lane];
    Key ak = {__shfl_sync(FULL, acc.d, k - 1), __shfl_sync(FULL, acc.i, k - 1)};
    for (int w = 1; w < WARPS; ++w) {
      const Key c = lists[w][warp][lane];
      unsigned m = __ballot_sync(FULL, lane < k && keyLess(c, ak));
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const Key x = {__shfl_sync(FULL, c.d, src), __shfl_sync(FULL, c.i, src)};
        if (keyLess(x, ak)) topkInsert(acc, ak, x, k, lane);
      }
    }
    if (lane < k) partial[((size_t)blockIdx.x * nqueries + q0 + warp) * k + lane] = acc;
  }
}

// one warp per query: merge the tiles' sorted lists, write the k results nearest first
__global__ void knnMergeTopkKernel(const Key* __restrict__ partial, int ntiles, int nqueries, int k, long long* __restrict__ idx,
                                   double* __restrict__ dist) {
  const unsigned FULL = 0xffffffffu;
  const int q = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (q >= nqueries) return;
  Key acc = {CUDART_INF, LLONG_MAX}, ak = acc;
  // the tiles' lists are one flat stream of ntiles * k keys: 32 keys per trip, 4 trips of loads in flight
  const int total = ntiles * k;
  for (int f0 = 0; f0 < total; f0 += 128) {
    Key c[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int f = f0 + u * 32 + lane;
      c[u].d = CUDART_INF; c[u].i = LLONG_MAX;
      if (f < total) c[u] = partial[((size_t)(f / k) * nqueries + q) * k + (f % k)];
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      unsigned m = __ballot_sync(FULL, keyLess(c[u], ak));
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const Key x = {__shfl_sync(FULL, c[u].d, src), __shfl_sync(FULL, c[u].i, src)};
        if (keyLess(x, ak)) topkInsert(acc, ak, x, k, lane);
      }
    }
  }
  if (lane < k) {
    idx[(size_t)q * k + lane] = acc.i != LLONG_MAX ? acc.i : -1;
    dist[(size_t)q * k + lane] = acc.d;
  }
}

// withinBall: strict d < radius, output in scan order (basic.cc:170-181)
__global__ void __launch_bounds__(KNN_THREADS)
withinBallKernel(ModelView mv, const double* __restrict__ nodes, const int32_t* __restrict__ cc, size_t n,
                 const double* __restrict__ queries, double radius, int32_t ccFilter, size_t cap,
                 long long* __restrict__ idx, long long* __restrict__ count) {
  const Tables T = makeTables(mv.blob, mv);
  __shared__ unsigned warpCnt[KNN_THREADS / 32];
  __shared__ unsigned long long baseS;
  const double* q = queries + (size_t)blockIdx.x * mv.nq;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) baseS = 0;
  __syncthreads();
  for (size_t start = 0; start < n; start += KNN_THREADS) {
    const size_t i = start + threadIdx.x;
    bool in = false;
    if (i < n && (ccFilter < 0 || cc[i] == ccFilter)) in = weighedDistanceDev(T, mv, nodes + i * (size_t)mv.nq, q) < radius;
    const unsigned m = __ballot_sync(0xffffffffu, in);
    if (lane == 0) warpCnt[warp] = __popc(m);
    __syncthreads();
    unsigned before = 0, total = 0;
    for (int w = 0; w < KNN_THREADS / 32; ++w) { if (w < warp) before += warpCnt[w]; total += warpCnt[w]; }
    const unsigned long long o = baseS + before + __popc(m & ((1u << lane) - 1u));
    if (in && o < cap) idx[(size_t)blockIdx.x * cap + o] = (long long)i;
    __syncthreads();
    if (threadIdx.x == 0) baseS += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) count[blockIdx.x] = (long long)baseS;
}

__global__ void relabelKernel(int32_t* cc, size_t n, int32_t from, int32_t to) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && cc[i] == from) cc[i] = to;
}

// exact scan of the first nNodes nodes: k selection passes or the single-pass top-k (see the kernels above)
void knnExactRun(hppb200_knn* nn, const double* d_q, size_t nqueries, int k, int32_t cc, size_t nNodes, long long* d_i,
                 double* d_d) {
  const size_t nq = (size_t)nn->model->nq;
  const size_t smemBytes = 16 + nn->view.blob_bytes + ((size_t)KNN_QB * nq + (size_t)nn->model->njoints) * sizeof(double);
  const bool ff = nn->ff;
  auto tileKernel = ff ? knnTileKernel<true> : knnTileKernel<false>;
  auto topkKernel = ff ? knnTileTopkKernel<true> : knnTileTopkKernel<false>;
  const unsigned groups = (unsigned)((nqueries + KNN_QB - 1) / KNN_QB);
  size_t tiles = std::max<size_t>(1, std::min<size_t>((size_t)nn->num_sms * 4, (nNodes + KNN_THREADS - 1) / KNN_THREADS));
  if (groups > 64) tiles = std::max<size_t>(1, std::min<size_t>(tiles, (size_t)nn->num_sms / 2));  // many groups fill the GPU anyway
  if (k > 1 && k <= 32) {  // single pass (k = 1 is one selection pass anyway, without the insertion logic)
    nn->d_keys.reserve(tiles * nqueries * (size_t)k * sizeof(Key) + 256);
    Key* d_part = (Key*)nn->d_keys.p;
    topkKernel<<<dim3((unsigned)tiles, groups), KNN_THREADS, smemBytes, nn->stream>>>(nn->view, nn->d_nodes, nn->d_cc, nNodes, d_q,
                                                                                        (int)nqueries, cc, k, d_part);
    knnMergeTopkKernel<<<(unsigned)((nqueries + 3) / 4), 128, 0, nn->stream>>>(d_part, (int)tiles, (int)nqueries, k, d_i, d_d);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    return;
  }
  nn->d_keys.reserve((nqueries + tiles * nqueries) * sizeof(Key) + 256);
  Key* d_prev = (Key*)nn->d_keys.p;
  Key* d_part = d_prev + nqueries;
  for (int r = 0; r < k; ++r) {
    tileKernel<<<dim3((unsigned)tiles, groups), KNN_THREADS, smemBytes, nn->stream>>>(nn->view, nn->d_nodes, nn->d_cc, nNodes, d_q,
                                                                                    (int)nqueries, cc, r ? d_prev : nullptr, d_part);
    knnReduceKernel<<<(unsigned)((nqueries + 3) / 4), 128, 0, nn->stream>>>(d_part, (int)tiles, (int)nqueries, k, r, d_prev, d_i, d_d);
  }
  HPPB200_CUDA_CHECK(cudaGetLastError());
}

void setLastError(const std::string& s);

template <typename F>
static int guardedKnn(F&& f) {
  try { f(); return HPPB200_OK; }
  catch (const std::invalid_argument& e) { setLastError(e.what()); return HPPB200_ERR_INVALID_ARGUMENT; }
  catch (const std::exception& e) { setLastError(e.what()); return std::strncmp(e.what(), "CUDA", 4) == 0 ? HPPB200_ERR_CUDA : HPPB200_ERR_RUNTIME; }
}

}  // namespace hppb200

using namespace hppb200;

extern "C" {

int hppb200_knn_create(const hppb200_model* m, int device, hppb200_knn** out) {
  return guardedKnn([&] {
    if (!m || !out) throw std::invalid_argument("null argument");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) throw std::runtime_error("CUDA: no device visible (no CPU fallback)");
    if (device < 0 || device >= count) throw std::invalid_argument("device index out of range");
    auto nn = new hppb200_knn;
    nn->model = &m->impl;
    nn->device = device;
    HPPB200_CUDA_CHECK(cudaSetDevice(device));
    nn->view = nn->model->deviceView(device);
    HPPB200_CUDA_CHECK(cudaStreamCreateWithFlags(&nn->stream, cudaStreamNonBlocking));
    HPPB200_CUDA_CHECK(cudaDeviceGetAttribute(&nn->num_sms, cudaDevAttrMultiProcessorCount, device));
    for (const DevJoint& J : nn->model->joints) nn->ff = nn->ff || J.type == HPPB200_JOINT_FREEFLYER;
    {
      const size_t smemBytes = 16 + nn->view.blob_bytes + ((size_t)KNN_QB * nn->model->nq + (size_t)nn->model->njoints) * sizeof(double);
      HPPB200_CUDA_CHECK(cudaFuncSetAttribute(nn->ff ? knnTileKernel<true> : knnTileKernel<false>,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
      HPPB200_CUDA_CHECK(cudaFuncSetAttribute(nn->ff ? knnTileTopkKernel<true> : knnTileTopkKernel<false>,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
      // pinned staging for small calls (a planner's single queries): truly asynchronous copies, one sync per call
      HPPB200_CUDA_CHECK(cudaHostAlloc((void**)&nn->h_stage, KNN_STAGE_BYTES, cudaHostAllocMapped));
      HPPB200_CUDA_CHECK(cudaHostGetDevicePointer((void**)&nn->d_stage_map, nn->h_stage, 0));  // the single-query kernel writes its result here
      if (const char* v = std::getenv("HPPB200_KNN_LATENCY")) nn->lat_enabled = std::atoi(v) != 0;
      if (const char* v = std::getenv("HPPB200_KNN_TIMING")) nn->lat_timing = std::atoi(v) != 0;
    }
    knnTcInit(nn);
    *out = nn;
  });
}
void hppb200_knn_destroy(hppb200_knn* nn) {
  if (!nn) return;
  cudaSetDevice(nn->device);
  if (nn->d_nodes) cudaFree(nn->d_nodes);
  if (nn->d_cc) cudaFree(nn->d_cc);
  knnTcDestroy(nn);
  if (nn->lat_ev0) cudaEventDestroy(nn->lat_ev0);
  if (nn->lat_ev1) cudaEventDestroy(nn->lat_ev1);
  if (nn->h_stage) cudaFreeHost(nn->h_stage);
  if (nn->stream) cudaStreamDestroy(nn->stream);
  delete nn;
}
size_t hppb200_knn_size(const hppb200_knn* nn) { return nn->n; }

int hppb200_knn_append(hppb200_knn* nn, const double* q, const int32_t* cc, size_t n) {
  return guardedKnn([&] {
    if (!nn || (n && !q)) throw std::invalid_argument("null argument");
    HPPB200_CUDA_CHECK(cudaSetDevice(nn->device));
    const size_t nq = (size_t)nn->model->nq;
    if (nn->n + n > nn->cap) {  // amortised growth, device-to-device move
      size_t ncap = std::max<size_t>((nn->n + n) * 2, 1024);
      double* nd = nullptr;
      int32_t* nc = nullptr;
      HPPB200_CUDA_CHECK(cudaMalloc((void**)&nd, ncap * nq * sizeof(double)));
      HPPB200_CUDA_CHECK(cudaMalloc((void**)&nc, ncap * sizeof(int32_t)));
      if (nn->n) {
        HPPB200_CUDA_CHECK(cudaMemcpyAsync(nd, nn->d_nodes, nn->n * nq * sizeof(double), cudaMemcpyDeviceToDevice, nn->stream));
        HPPB200_CUDA_CHECK(cudaMemcpyAsync(nc, nn->d_cc, nn->n * sizeof(int32_t), cudaMemcpyDeviceToDevice, nn->stream));
        HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
      }
      if (nn->d_nodes) cudaFree(nn->d_nodes);
      if (nn->d_cc) cudaFree(nn->d_cc);
      nn->d_nodes = nd; nn->d_cc = nc; nn->cap = ncap;
    }
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(nn->d_nodes + nn->n * nq, q, n * nq * sizeof(double), cudaMemcpyHostToDevice, nn->stream));
    if (cc) HPPB200_CUDA_CHECK(cudaMemcpyAsync(nn->d_cc + nn->n, cc, n * sizeof(int32_t), cudaMemcpyHostToDevice, nn->stream));
    else HPPB200_CUDA_CHECK(cudaMemsetAsync(nn->d_cc + nn->n, 0, n * sizeof(int32_t), nn->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
    nn->n += n;
  });
}

int hppb200_knn_relabel(hppb200_knn* nn, int32_t from_cc, int32_t to_cc) {
  return guardedKnn([&] {
    HPPB200_CUDA_CHECK(cudaSetDevice(nn->device));
    if (nn->n == 0) return;
    relabelKernel<<<(unsigned)((nn->n + 255) / 256), 256, 0, nn->stream>>>(nn->d_cc, nn->n, from_cc, to_cc);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
  });
}

int hppb200_knn_set_labels(hppb200_knn* nn, const int32_t* cc, size_t n) {
  return guardedKnn([&] {
    if (!nn || (n && !cc)) throw std::invalid_argument("null argument");
    if (n != nn->n) throw std::invalid_argument("label count differs from the number of nodes");
    if (n == 0) return;
    HPPB200_CUDA_CHECK(cudaSetDevice(nn->device));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(nn->d_cc, cc, n * sizeof(int32_t), cudaMemcpyHostToDevice, nn->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
  });
}

int hppb200_knn_query(hppb200_knn* nn, const double* queries, size_t nqueries, int k, int32_t cc, uint32_t flags,
                      int64_t* idx, double* dist) {
  return guardedKnn([&] {
    if (!nn || (nqueries && (!queries || !idx || !dist)) || k <= 0) throw std::invalid_argument("bad argument");
    if (nqueries == 0) return;
    HPPB200_CUDA_CHECK(cudaSetDevice(nn->device));
    const size_t nq = (size_t)nn->model->nq;
    if (nqueries > 0x7fffffffull / 4) throw std::invalid_argument("too many queries in one call");
    auto run = [&](const double* d_q, long long* d_i, double* d_d) {
      nn->tc_last_used = false;
      if (knnTcEligible(nn, nqueries, k, cc)) knnTcRun(nn, d_q, nqueries, k, cc, d_i, d_d);
      else knnExactRun(nn, d_q, nqueries, k, cc, nn->n, d_i, d_d);
    };
    if (!(flags & HPPB200_DEVICE_POINTERS) && knnLatencyEligible(nn, nqueries, k)) {
      // a planner's single query: one kernel (query in the parameter block, result written to mapped pinned memory by
      // the last CTA), one synchronisation
      nn->tc_last_used = false;
      nn->lat_last_used = true;
      knnLatencyRun(nn, queries, k, cc, (long long*)nn->d_stage_map, (double*)(nn->d_stage_map + 32 * sizeof(long long)));
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
      std::memcpy(idx, nn->h_stage, (size_t)k * sizeof(int64_t));
      std::memcpy(dist, nn->h_stage + 32 * sizeof(long long), (size_t)k * sizeof(double));
      return;
    }
    nn->lat_last_used = false;
    if (flags & HPPB200_DEVICE_POINTERS) {
      run(queries, (long long*)idx, dist);
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
      return;
    }
    nn->d_queries.reserve(nqueries * nq);
    nn->d_idx.reserve(nqueries * k);
    nn->d_dist.reserve(nqueries * k);
    const size_t qBytes = nqueries * nq * sizeof(double), rBytes = nqueries * (size_t)k * sizeof(double);
    if (qBytes + 2 * rBytes <= KNN_STAGE_BYTES) {  // small call: through the pinned staging block
      unsigned char* h = nn->h_stage;
      std::memcpy(h, queries, qBytes);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(nn->d_queries.p, h, qBytes, cudaMemcpyHostToDevice, nn->stream));
      run(nn->d_queries.p, nn->d_idx.p, nn->d_dist.p);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(h + qBytes, nn->d_idx.p, rBytes, cudaMemcpyDeviceToHost, nn->stream));
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(h + qBytes + rBytes, nn->d_dist.p, rBytes, cudaMemcpyDeviceToHost, nn->stream));
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
      std::memcpy(idx, h + qBytes, rBytes);
      std::memcpy(dist, h + qBytes + rBytes, rBytes);
      return;
    }
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(nn->d_queries.p, queries, qBytes, cudaMemcpyHostToDevice, nn->stream));
    run(nn->d_queries.p, nn->d_idx.p, nn->d_dist.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(idx, nn->d_idx.p, rBytes, cudaMemcpyDeviceToHost, nn->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(dist, nn->d_dist.p, rBytes, cudaMemcpyDeviceToHost, nn->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
  });
}

int hppb200_knn_within_ball(hppb200_knn* nn, const double* queries, size_t nqueries, double radius, int32_t cc,
                            size_t cap, int64_t* idx, int64_t* count) {
  return guardedKnn([&] {
    if (!nn || (nqueries && (!queries || !idx || !count))) throw std::invalid_argument("bad argument");
    if (nqueries == 0) return;
    HPPB200_CUDA_CHECK(cudaSetDevice(nn->device));
    const size_t nq = (size_t)nn->model->nq;
    nn->d_queries.reserve(nqueries * nq);
    nn->d_idx.reserve(nqueries * cap);
    nn->d_count.reserve(nqueries);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(nn->d_queries.p, queries, nqueries * nq * sizeof(double), cudaMemcpyHostToDevice, nn->stream));
    withinBallKernel<<<(unsigned)nqueries, KNN_THREADS, 0, nn->stream>>>(nn->view, nn->d_nodes, nn->d_cc, nn->n, nn->d_queries.p, radius, cc,
                                                                       cap, nn->d_idx.p, nn->d_count.p);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(idx, nn->d_idx.p, nqueries * cap * sizeof(int64_t), cudaMemcpyDeviceToHost, nn->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(count, nn->d_count.p, nqueries * sizeof(int64_t), cudaMemcpyDeviceToHost, nn->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
  });
}

}  // extern "C"
