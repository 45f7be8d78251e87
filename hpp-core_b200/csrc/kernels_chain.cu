// PathVector recursion of ContinuousValidation::validate (reference src/continuous-validation.cc:153-192) for chains of
// StraightPath leaves.  The reference validates the sub-paths one after the other and stops at the first invalid one;
// the sub-path validations are independent, so the engine validates ALL leaves of ALL chains as one edge batch
// (hppb200_validate_edges machinery) and this kernel replays the recursion's bookkeeping per chain: which leaf fails
// first in walking order, the parameter offset added to its report (`param`, accumulated leaf by leaf exactly like
// the reference: += length() going forward, -= length() going backward from PathVector::length()), where the valid
// part ends.  Results of leaves behind the first failure are simply not looked at.
#include <math_constants.h>

#include "engine.hpp"

namespace hppb200 {

__global__ void chainReduceKernel(const int32_t* __restrict__ chainBegin, size_t nChains, int reverse, const double* __restrict__ L,
                                  const uint8_t* __restrict__ eValid, const double* __restrict__ eTfail,
                                  const double* __restrict__ eTlast, const int32_t* __restrict__ ePair,
                                  const uint8_t* __restrict__ eReason, uint8_t* valid, double* tFail, double* tLast, int32_t* pair,
                                  uint8_t* reason, int32_t* firstBad) {
  const size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nChains) return;
  const int b = chainBegin[c], e = chainBegin[c + 1];
  int bad = -1;
  double param = 0, tf = CUDART_NAN, tl = 0;
  if (!reverse) {
    for (int i = b; i < e; ++i) {
      if (eValid[i]) param += L[i];  // :183-185
      else { bad = i; tf = eTfail[i] + param; tl = param + eTlast[i]; break; }  // :186-189
    }
    if (bad < 0) tl = param;
  } else {
    for (int i = b; i < e; ++i) param += L[i];  // PathVector::length(): timeRange accumulated by appendPath
    for (int i = e - 1; i >= b; --i) {
      if (eValid[i]) param -= L[i];  // :163-165
      else { bad = i; tf = eTfail[i] + (param - L[i]); tl = (param - L[i]) + eTlast[i]; break; }  // :166-174
    }
    if (bad < 0) tl = 0.0;
  }
  valid[c] = bad < 0 ? 1 : 0;
  if (tFail) tFail[c] = tf;
  if (tLast) tLast[c] = tl;
  if (pair) pair[c] = bad < 0 ? -1 : ePair[bad];
  if (reason) reason[c] = bad < 0 ? (uint8_t)HPPB200_REASON_VALID : eReason[bad];
  if (firstBad) firstBad[c] = bad;
}

void launchChainReduce(Workspace* ws, const int32_t* d_chainBegin, size_t nChains, int reverse, const double* d_L, const uint8_t* eValid,
                       const double* eTfail, const double* eTlast, const int32_t* ePair, const uint8_t* eReason, uint8_t* valid,
                       double* tFail, double* tLast, int32_t* pair, uint8_t* reason, int32_t* firstBad) {
  if (!nChains) return;
  chainReduceKernel<<<(unsigned)((nChains + 127) / 128), 128, 0, ws->stream>>>(d_chainBegin, nChains, reverse, d_L, eValid, eTfail, eTlast,
                                                                              ePair, eReason, valid, tFail, tLast, pair, reason, firstBad);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
}

}  // namespace hppb200

// ------------------------------------------------------------------------------------------------
// Exchange helpers of the sharded (one process per GPU) form of the path, SURVEY.md 8e: the only data that crosses
// GPUs are validity masks and k-NN candidates.
namespace hppb200 {

// valid [n] u8 (0 / 1) -> bits [(n + 31) / 32] u32, bit i % 32 of word i / 32: the all-gather then moves n / 8 bytes
// per rank instead of n (1 MiB -> 128 KiB per 1 M configurations; the collective is latency-bound).
__global__ void packMaskKernel(const uint8_t* __restrict__ valid, size_t n, uint32_t* __restrict__ bits) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned b = __ballot_sync(0xffffffffu, i < n && valid[i] != 0);
  if ((threadIdx.x & 31) == 0 && i < n) bits[i >> 5] = b;
}
__global__ void unpackMaskKernel(const uint32_t* __restrict__ bits, size_t n, uint8_t* __restrict__ valid) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) valid[i] = (bits[i >> 5] >> (i & 31)) & 1u;
}

// W-way merge of per-shard k-NN candidates, one thread per query.  dist / idx: [W][Q][k] (rank-major, what an
// all-gather of the per-rank [Q][k] results produces), every list nearest first by the key (distance, GLOBAL node
// index), padded with idx = -1.  out: the k smallest keys over all shards, same order and padding: the reference's
// tie rule (strict '<' keeps the earlier node, src/nearest-neighbor/basic.cc:66,106) carried across shards.
__global__ void knnMergeShardsKernel(const double* __restrict__ dist, const long long* __restrict__ idx, int W, size_t Q, int k,
                                     long long* __restrict__ outIdx, double* __restrict__ outDist) {
  const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  int head[16];
  for (int w = 0; w < W; ++w) head[w] = 0;
  for (int j = 0; j < k; ++j) {
    int best = -1;
    double bd = 0;
    long long bi = 0;
    for (int w = 0; w < W; ++w) {
      if (head[w] >= k) continue;
      const size_t o = ((size_t)w * Q + q) * k + head[w];
      const long long ci = idx[o];
      if (ci < 0) { head[w] = k; continue; }  // padding: this shard is exhausted
      const double cd = dist[o];
      if (best < 0 || cd < bd || (cd == bd && ci < bi)) { best = w; bd = cd; bi = ci; }
    }
    if (best < 0) { outIdx[q * k + j] = -1; outDist[q * k + j] = CUDART_INF; }
    else { outIdx[q * k + j] = bi; outDist[q * k + j] = bd; ++head[best]; }
  }
}

void launchPackMask(cudaStream_t stream, const uint8_t* valid, size_t n, uint32_t* bits) {
  if (!n) return;
  packMaskKernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(valid, n, bits);
  HPPB200_CUDA_CHECK(cudaGetLastError());
}
void launchUnpackMask(cudaStream_t stream, const uint32_t* bits, size_t n, uint8_t* valid) {
  if (!n) return;
  unpackMaskKernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(bits, n, valid);
  HPPB200_CUDA_CHECK(cudaGetLastError());
}
void launchKnnMergeShards(cudaStream_t stream, const double* dist, const int64_t* idx, int W, size_t Q, int k, int64_t* outIdx, double* outDist) {
  if (!Q || k <= 0) return;
  if (W < 1 || W > 16) throw std::invalid_argument("knn shard merge: 1..16 shards");
  knnMergeShardsKernel<<<(unsigned)((Q + 127) / 128), 128, 0, stream>>>(dist, (const long long*)idx, W, Q, k, (long long*)outIdx, outDist);
  HPPB200_CUDA_CHECK(cudaGetLastError());
}

}  // namespace hppb200

// ------------------------------------------------------------------------------------------------
// Edges given as pairs of ROADMAP NODE indices (kPRM*-style linking, src/path-planner/k-prm-star.cc:134-193: every
// candidate edge joins two nodes that already live in the device mirror of the roadmap): the end configurations are
// gathered on the device instead of being gathered on the host and uploaded (96 B per UR5 edge down to 16 B).
namespace hppb200 {
__global__ void gatherEdgeEndsKernel(const double* __restrict__ nodes, size_t nNodes, int nq, const long long* __restrict__ a,
                                     const long long* __restrict__ b, size_t n, double* __restrict__ q1, double* __restrict__ q2,
                                     unsigned* __restrict__ bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * (size_t)nq) return;
  const size_t e = i / (size_t)nq;
  const int c = (int)(i - e * (size_t)nq);
  const long long ia = a[e], ib = b[e];
  if (ia < 0 || ib < 0 || (size_t)ia >= nNodes || (size_t)ib >= nNodes) { if (c == 0) atomicAdd(bad, 1u); q1[i] = 0; q2[i] = 0; return; }
  q1[i] = nodes[(size_t)ia * nq + c];
  q2[i] = nodes[(size_t)ib * nq + c];
}
void launchGatherEdgeEnds(cudaStream_t stream, const double* nodes, size_t nNodes, int nq, const int64_t* a, const int64_t* b, size_t n,
                          double* q1, double* q2, unsigned* bad) {
  if (!n) return;
  const size_t total = n * (size_t)nq;
  gatherEdgeEndsKernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(nodes, nNodes, nq, (const long long*)a, (const long long*)b, n, q1, q2, bad);
  HPPB200_CUDA_CHECK(cudaGetLastError());
}
}  // namespace hppb200
