// First-sample screen of batched continuous edge validation (Progressive and Dichotomy on straight paths).
//
// Both methods start by validating the configuration at the path's first parameter t0 (0, or L in reverse:
// progressive.cc:93-100, dichotomy.cc:84-90), with no validated interval on any body pair yet, so
// ContinuousValidation::validateConfiguration (continuous-validation.cc:128-148) walks ALL body pairs in construction
// order and BodyPairCollision::computeDistanceLowerBound (body-pair-collision.cc:175-198) stops at the first colliding
// geometry pair: the edge fails with t_fail = t_last_valid = t0 and that pair.  That is a plain configuration
// validation against the continuous pair table, which the configuration kernels (kernels_config.cu: thread per
// configuration, exact culls, GJK job queue) do two orders of magnitude faster per sample than an edge engine
// (CTA per edge, or a slot of the wavefront).  So a batch is screened first:
//   1. q(t0) of every edge (StraightPath::impl_compute returns the stored end configurations at the range ends,
//      straight-path.cc:91-110: q1, or q2 in reverse) through configPassKernel + gjkJobsKernel on the body pairs'
//      geometry pairs (launchFirstSampleScreen);
//   2. edges whose first sample collides get their final results here; the survivors are compacted (stable order),
//      gathered into a contiguous sub-batch, validated by the edge engines from scratch, and scattered back.
// Results are those of the unscreened path bit for bit (tests/test_gpu_parity.py::test_first_sample_screen_*).
// Batches whose start configurations are all valid (a planner extending from roadmap nodes) pay the screen (~1 % of a
// Progressive batch) and run unscreened; batches of random edges lose the share that starts in collision (93 % on
// BASELINE configs[3]).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>

#include <chrono>
#include <cstdio>
#include <cstdlib>

#include "engine.hpp"

namespace hppb200 {

void launchValidateEdgesCore(Workspace* ws, int method, double param, int reverse, const double* d_q1, const double* d_q2,
                             const double* d_L, size_t n, uint32_t flags, uint8_t* d_valid, double* d_tfail, double* d_tlast,
                             int32_t* d_pair, uint8_t* d_reason);

namespace {

// reverse: q(t0) = q(L) = q2, except on a zero-length path (t0 = 0: q1)
__global__ void fsStartConfigKernel(const double* __restrict__ q1, const double* __restrict__ q2, const double* __restrict__ L, size_t n,
                                    int nq, double* __restrict__ q) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n * (size_t)nq) return;
  const size_t i = k / (size_t)nq;
  q[k] = L[i] == 0.0 ? q1[k] : q2[k];
}

// final results of the edges whose first sample collides (what every edge engine writes for them)
__global__ void fsFinishKernel(const uint8_t* __restrict__ fsValid, const int32_t* __restrict__ fsFirst, const double* __restrict__ L,
                               int reverse, size_t n, uint8_t* valid, double* tfail, double* tlast, int32_t* pair, uint8_t* reason,
                               DeviceStats* stats) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool hit = i < n && fsValid[i] == 0;
  if (hit) {
    const double t0 = reverse ? L[i] : 0.0;
    valid[i] = 0;
    if (tfail) tfail[i] = t0;
    if (tlast) tlast[i] = t0;
    if (pair) pair[i] = fsFirst[i];
    if (reason) reason[i] = HPPB200_REASON_COLLISION;
  }
  const unsigned b = __ballot_sync(0xffffffffu, hit);
  if ((threadIdx.x & 31) == 0 && b && stats) atomicAdd(&stats->edge_samples, (unsigned long long)__popc(b));
}

__global__ void fsKeysKernel(const uint32_t* __restrict__ list, size_t ns, const unsigned long long* __restrict__ dmin, unsigned long long* __restrict__ keys) {
  const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j < ns) keys[j] = dmin[list[j]];
}

__global__ void fsGatherKernel(const uint32_t* __restrict__ list, size_t ns, int nq, const double* __restrict__ q1,
                               const double* __restrict__ q2, const double* __restrict__ L, double* __restrict__ q1s,
                               double* __restrict__ q2s, double* __restrict__ Ls) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ns * (size_t)nq) return;
  const size_t j = k / (size_t)nq, c = k - j * (size_t)nq;
  const size_t src = (size_t)list[j] * (size_t)nq + c;
  q1s[k] = q1[src];
  q2s[k] = q2[src];
  if (c == 0) Ls[j] = L[list[j]];
}

__global__ void fsScatterKernel(const uint32_t* __restrict__ list, size_t ns, const uint8_t* __restrict__ v, const double* __restrict__ tf,
                                const double* __restrict__ tl, const int32_t* __restrict__ pr, const uint8_t* __restrict__ rs, uint8_t* valid,
                                double* tfail, double* tlast, int32_t* pair, uint8_t* reason) {
  const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= ns) return;
  const size_t i = list[j];
  valid[i] = v[j];
  if (tfail) tfail[i] = tf[j];
  if (tlast) tlast[i] = tl[j];
  if (pair) pair[i] = pr[j];
  if (reason) reason[i] = rs[j];
}

}  // namespace

void launchValidateEdges(Workspace* ws, int method, double param, int reverse, const double* d_q1, const double* d_q2, const double* d_L,
                         size_t n, uint32_t flags, uint8_t* d_valid, double* d_tfail, double* d_tlast, int32_t* d_pair, uint8_t* d_reason) {
  if (n == 0) return;
  // host-pointer calls: q2 / L may still be on their way (capi.cu); waited for right before they are first read
  cudaEvent_t inputsReady = ws->edge_inputs_ready;
  ws->edge_inputs_ready = nullptr;
  auto waitInputs = [&]() {
    if (!inputsReady) return;
    HPPB200_CUDA_CHECK(cudaStreamWaitEvent(ws->stream, inputsReady, 0));
    inputsReady = nullptr;
  };
  const bool screen = ws->first_sample_screen && method != HPPB200_METHOD_DISCRETIZED && n >= ws->first_sample_min && n <= 0x7fffffffull &&
                      ws->max_iterations >= 1 && ws->view.nbgpairs > 0 && firstSampleScreenAvailable(ws);
  if (!screen) {
    waitInputs();
    launchValidateEdgesCore(ws, method, param, reverse, d_q1, d_q2, d_L, n, flags, d_valid, d_tfail, d_tlast, d_pair, d_reason);
    return;
  }
  const bool trace = std::getenv("HPPB200_WF_TRACE") != nullptr;
  const int nq = ws->model->nq;
  const size_t chunk = std::min(n, firstSampleChunk()), chunks = (n + chunk - 1) / chunk;
  constexpr size_t kMaxChunks = 8192;  // (2^31 edges / 262144)
  if (!ws->h_fs) HPPB200_CUDA_CHECK(cudaHostAlloc((void**)&ws->h_fs, (1 + kMaxChunks) * sizeof(unsigned long long), cudaHostAllocDefault));
  const double* qStart = d_q1;
  if (reverse) {
    waitInputs();
    ws->d_fs_q.reserve(n * (size_t)nq);
    fsStartConfigKernel<<<(unsigned)((n * (size_t)nq + 255) / 256), 256, 0, ws->stream>>>(d_q1, d_q2, d_L, n, nq, ws->d_fs_q.p);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    ws->stats.kernel_launches += 1;
    qStart = ws->d_fs_q.p;
  }
  ws->d_fs_valid.reserve(n);
  ws->d_fs_first.reserve(n);
  ws->d_fs_reason.reserve(n);
  ws->d_fs_cnt.reserve(1 + chunks);
  ws->d_fs_list.reserve(n);
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (trace) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, ws->stream); }
  const bool order = ws->first_sample_order;
  if (order) ws->d_fs_dmin.reserve(n);
  const bool hint = launchFirstSampleScreen(ws, qStart, n, ws->d_fs_valid.p, ws->d_fs_first.p, ws->d_fs_reason.p, ws->d_fs_cnt.p + 1,
                                            order ? ws->d_fs_dmin.p : nullptr);
  if (trace) cudaEventRecord(e1, ws->stream);
  // survivors, in input order
  size_t tmpBytes = 0;
  cub::CountingInputIterator<uint32_t> iota(0u);
  cub::DeviceSelect::Flagged(nullptr, tmpBytes, iota, ws->d_fs_valid.p, ws->d_fs_list.p, ws->d_fs_cnt.p, (int)n, ws->stream);
  ws->d_fs_tmp.reserve(tmpBytes + 256);
  HPPB200_CUDA_CHECK(cub::DeviceSelect::Flagged(ws->d_fs_tmp.p, tmpBytes, iota, ws->d_fs_valid.p, ws->d_fs_list.p, ws->d_fs_cnt.p, (int)n, ws->stream));
  ws->stats.kernel_launches += 2;
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->h_fs, ws->d_fs_cnt.p, (1 + chunks) * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ws->stream));
  waitInputs();  // (everything below reads q2 / L; the synchronisation that follows covers their copies as well)
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  const size_t ns = (size_t)ws->h_fs[0];
  bool overflow = false;
  for (size_t c = 0; c < chunks; ++c) overflow |= ws->h_fs[1 + c] > firstSampleQueueCap(chunk);
  if (trace) {
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    unsigned long long jobs = 0;
    for (size_t c = 0; c < chunks; ++c) jobs += ws->h_fs[1 + c];
    std::fprintf(stderr, "[edges] first-sample screen: %zu of %zu edges go on%s; configuration kernels %.3f ms, %.2f GJK jobs per edge\n", ns, n,
                 overflow ? " (job queue overflow: screen dropped)" : "", ms, (double)jobs / (double)n);
  }
  // a screen that removes (almost) nothing: the engines take the batch as it is
  if (overflow || ns > n || ns * 10 > n * 9) {
    launchValidateEdgesCore(ws, method, param, reverse, d_q1, d_q2, d_L, n, flags, d_valid, d_tfail, d_tlast, d_pair, d_reason);
    return;
  }
  fsFinishKernel<<<(unsigned)((n + 255) / 256), 256, 0, ws->stream>>>(ws->d_fs_valid.p, ws->d_fs_first.p, d_L, reverse, n, d_valid, d_tfail,
                                                                     d_tlast, d_pair, d_reason, ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  if (ns == 0) return;
  ws->d_fs_q1.reserve(ns * (size_t)nq);
  ws->d_fs_q2.reserve(ns * (size_t)nq);
  ws->d_fs_L.reserve(ns);
  ws->d_fs_ovalid.reserve(ns);
  ws->d_fs_tfail.reserve(ns);
  ws->d_fs_tlast.reserve(ns);
  ws->d_fs_opair.reserve(ns);
  ws->d_fs_oreason.reserve(ns);
  // The edges that go on, closest-to-an-obstacle first: an edge that starts (or, in reverse, ends) millimetres from an obstacle is
  // the one that takes hundreds of dependent samples, and the CTA-per-edge kernel ends when the LAST such chain ends: started in
  // input order, the longest chains of BASELINE configs[3] begin anywhere in the batch and stretch it by ~20 ms of 63.  The order
  // decides only when an edge runs; the key is the smallest distance the screen happened to evaluate at q(t0).
  const uint32_t* list = ws->d_fs_list.p;
  if (hint) {
    ws->d_fs_keys.reserve(ns);
    ws->d_fs_keys2.reserve(ns);
    ws->d_fs_list2.reserve(ns);
    fsKeysKernel<<<(unsigned)((ns + 255) / 256), 256, 0, ws->stream>>>(ws->d_fs_list.p, ns, ws->d_fs_dmin.p, ws->d_fs_keys.p);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    size_t sortBytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sortBytes, ws->d_fs_keys.p, ws->d_fs_keys2.p, ws->d_fs_list.p, ws->d_fs_list2.p, (int)ns, 0, 64, ws->stream);
    ws->d_fs_tmp.reserve(sortBytes + 256);
    HPPB200_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(ws->d_fs_tmp.p, sortBytes, ws->d_fs_keys.p, ws->d_fs_keys2.p, ws->d_fs_list.p, ws->d_fs_list2.p,
                                                       (int)ns, 0, 64, ws->stream));
    ws->stats.kernel_launches += 2;
    list = ws->d_fs_list2.p;
  }
  fsGatherKernel<<<(unsigned)((ns * (size_t)nq + 255) / 256), 256, 0, ws->stream>>>(list, ns, nq, d_q1, d_q2, d_L, ws->d_fs_q1.p,
                                                                                   ws->d_fs_q2.p, ws->d_fs_L.p);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  std::chrono::steady_clock::time_point c0;
  if (trace) { cudaStreamSynchronize(ws->stream); c0 = std::chrono::steady_clock::now(); }
  launchValidateEdgesCore(ws, method, param, reverse, ws->d_fs_q1.p, ws->d_fs_q2.p, ws->d_fs_L.p, ns, flags, ws->d_fs_ovalid.p,
                          ws->d_fs_tfail.p, ws->d_fs_tlast.p, ws->d_fs_opair.p, ws->d_fs_oreason.p);
  if (trace) {
    cudaStreamSynchronize(ws->stream);
    std::fprintf(stderr, "[edges] engines on the %zu screened edges: %.3f ms\n", ns,
                 std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - c0).count());
  }
  fsScatterKernel<<<(unsigned)((ns + 255) / 256), 256, 0, ws->stream>>>(list, ns, ws->d_fs_ovalid.p, ws->d_fs_tfail.p, ws->d_fs_tlast.p,
                                                                       ws->d_fs_opair.p, ws->d_fs_oreason.p, d_valid, d_tfail, d_tlast, d_pair,
                                                                       d_reason);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 2;
}

}  // namespace hppb200
