// Brute-force WeighedDistance nearest neighbour over roadmap nodes mirrored on the device
// (reference src/nearest-neighbor/basic.cc:52-181, src/weighed-distance.cc:272-289).
//
// Exact FP64 path: one CTA per query scans all nodes (coalesced, 8*nq bytes per node: the
// kernel is HBM/L2-bound on the node stream) and selects the k smallest by the
// lexicographic key (distance, node index), which is exactly what the reference's linear
// scan with strict '<' keeps: on equal distances the earlier node wins
// (basic.cc:66,106).  k-NN is k selection passes (k is small: 1 for BiRRT, O(log N) for
// kPRM*); every pass keeps the best key strictly greater than the previous one.
#include <climits>
#include <cstring>

#include "dev_model.cuh"

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
};

namespace hppb200 {

constexpr int KNN_THREADS = 256;

__device__ __forceinline__ double weighedDistanceDev(const Tables& T, const ModelView& mv, const double* __restrict__ a,
                                                     const double* __restrict__ b) {
  double res = 0;
  for (int j = 0; j < mv.njoints; ++j) {
    const DevJoint& J = T.joints[j];
    const double w = T.weights[j];
    const double w2 = w * w;
    const double dq = b[J.idx_q] - a[J.idx_q];
    res += (J.type <= HPPB200_JOINT_REVOLUTE_Z) ? w2 * (dq * dq) : dq * dq;
  }
  return sqrt(res);
}

struct Key {
  double d;
  long long i;
};
__device__ __forceinline__ bool keyLess(const Key& a, const Key& b) { return a.d < b.d || (a.d == b.d && a.i < b.i); }

__global__ void __launch_bounds__(KNN_THREADS)
knnQueryKernel(ModelView mv, const double* __restrict__ nodes, const int32_t* __restrict__ cc, size_t n,
               const double* __restrict__ queries, int k, int32_t ccFilter, long long* __restrict__ idx,
               double* __restrict__ dist) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  unsigned char* sblob = smem + 16;
  stageBlobTMA(sblob, mv.blob, mv.blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);
  __shared__ Key red[KNN_THREADS / 32];
  __shared__ Key prevS;
  const double* q = queries + (size_t)blockIdx.x * mv.nq;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  Key prev = {-1.0, -1};
  for (int r = 0; r < k; ++r) {
    Key best = {CUDART_INF, LLONG_MAX};
    for (size_t i = threadIdx.x; i < n; i += KNN_THREADS) {
      if (ccFilter >= 0 && cc[i] != ccFilter) continue;
      Key c = {weighedDistanceDev(T, mv, nodes + i * (size_t)mv.nq, q), (long long)i};
      if (keyLess(prev, c) && keyLess(c, best)) best = c;
    }
    for (int o = 16; o > 0; o >>= 1) {
      Key other = {__shfl_xor_sync(0xffffffffu, best.d, o), __shfl_xor_sync(0xffffffffu, best.i, o)};
      if (keyLess(other, best)) best = other;
    }
    if (lane == 0) red[warp] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
      Key b = red[0];
      for (int w = 1; w < KNN_THREADS / 32; ++w)
        if (keyLess(red[w], b)) b = red[w];
      prevS = b;
      const bool found = b.i != LLONG_MAX;
      idx[(size_t)blockIdx.x * k + r] = found ? b.i : -1;
      dist[(size_t)blockIdx.x * k + r] = b.d;
    }
    __syncthreads();
    prev = prevS;
    if (prev.i == LLONG_MAX) {  // fewer than k candidates: pad the rest
      if (threadIdx.x == 0)
        for (int s = r + 1; s < k; ++s) { idx[(size_t)blockIdx.x * k + s] = -1; dist[(size_t)blockIdx.x * k + s] = CUDART_INF; }
      break;
    }
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
    *out = nn;
  });
}
void hppb200_knn_destroy(hppb200_knn* nn) {
  if (!nn) return;
  cudaSetDevice(nn->device);
  if (nn->d_nodes) cudaFree(nn->d_nodes);
  if (nn->d_cc) cudaFree(nn->d_cc);
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

int hppb200_knn_query(hppb200_knn* nn, const double* queries, size_t nqueries, int k, int32_t cc, uint32_t flags,
                      int64_t* idx, double* dist) {
  return guardedKnn([&] {
    if (!nn || (nqueries && (!queries || !idx || !dist)) || k <= 0) throw std::invalid_argument("bad argument");
    if (nqueries == 0) return;
    HPPB200_CUDA_CHECK(cudaSetDevice(nn->device));
    const size_t nq = (size_t)nn->model->nq;
    const size_t smemBytes = 16 + nn->view.blob_bytes;
    HPPB200_CUDA_CHECK(cudaFuncSetAttribute(knnQueryKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
    if (flags & HPPB200_DEVICE_POINTERS) {
      knnQueryKernel<<<(unsigned)nqueries, KNN_THREADS, smemBytes, nn->stream>>>(nn->view, nn->d_nodes, nn->d_cc, nn->n, queries, k, cc,
                                                                               (long long*)idx, dist);
      HPPB200_CUDA_CHECK(cudaGetLastError());
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
      return;
    }
    nn->d_queries.reserve(nqueries * nq);
    nn->d_idx.reserve(nqueries * k);
    nn->d_dist.reserve(nqueries * k);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(nn->d_queries.p, queries, nqueries * nq * sizeof(double), cudaMemcpyHostToDevice, nn->stream));
    knnQueryKernel<<<(unsigned)nqueries, KNN_THREADS, smemBytes, nn->stream>>>(nn->view, nn->d_nodes, nn->d_cc, nn->n, nn->d_queries.p, k,
                                                                             cc, nn->d_idx.p, nn->d_dist.p);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(idx, nn->d_idx.p, nqueries * k * sizeof(int64_t), cudaMemcpyDeviceToHost, nn->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(dist, nn->d_dist.p, nqueries * k * sizeof(double), cudaMemcpyDeviceToHost, nn->stream));
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
