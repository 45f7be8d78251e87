// Device pieces shared by the exact FP64 nearest-neighbour scan (kernels_knn.cu) and the exact re-rank of the
// tensor-core candidate pass (kernels_knn_tc.cu): WeighedDistance (reference src/weighed-distance.cc:272-289) in
// ONE operation order, the (distance, node index) key and the warp-register top-k list.
#pragma once
#include "dev_model.cuh"
#include "knn_state.hpp"

namespace hppb200 {

__device__ __forceinline__ double weighedDistanceDev(const Tables& T, const ModelView& mv, const double* __restrict__ a,
                                                     const double* __restrict__ b) {
  double res = 0;
  for (int j = 0; j < mv.njoints; ++j) {
    const DevJoint& J = T.joints[j];
    const double w = T.weights[j];
    const double w2 = w * w;
    if (J.type == HPPB200_JOINT_FREEFLYER) { res += ffSquaredDistance(a + J.idx_q, b + J.idx_q, w2); continue; }
    const double dq = b[J.idx_q] - a[J.idx_q];
    res += (J.type <= HPPB200_JOINT_REVOLUTE_Z) ? w2 * (dq * dq) : dq * dq;
  }
  return sqrt(res);
}

// WeighedDistance of one node to KNN_QB queries at once: the node entry, the joint record and w^2 are loaded once
// per joint; per query the operation sequence is the one of weighedDistanceDev (same bits).  `sq` holds the
// queries in shared memory, `sw2` the per-joint w * w.
template <int QB, bool FF>
__device__ __forceinline__ void weighedDistanceBatch(const Tables& T, const ModelView& mv, const double* __restrict__ node,
                                                     const double* sq, const double* sw2, double (&d)[QB]) {
#pragma unroll
  for (int q = 0; q < QB; ++q) d[q] = 0;
  for (int j = 0; j < mv.njoints; ++j) {
    const DevJoint& J = T.joints[j];
    const double w2 = sw2[j];
    if (FF && J.type == HPPB200_JOINT_FREEFLYER) {  // distance(node, query): the argument order of nearestNeighbor::Basic
#pragma unroll
      for (int q = 0; q < QB; ++q) d[q] += ffSquaredDistance(node + J.idx_q, sq + q * mv.nq + J.idx_q, w2);
      continue;
    }
    const double a = node[J.idx_q];
    const bool rev = J.type <= HPPB200_JOINT_REVOLUTE_Z;
#pragma unroll
    for (int q = 0; q < QB; ++q) {
      const double dq = sq[q * mv.nq + J.idx_q] - a;
      d[q] += rev ? w2 * (dq * dq) : dq * dq;
    }
  }
#pragma unroll
  for (int q = 0; q < QB; ++q) d[q] = sqrt(d[q]);
}

using Key = KnnKey;
__device__ __forceinline__ bool keyLess(const Key& a, const Key& b) { return a.d < b.d || (a.d == b.d && a.i < b.i); }

__device__ __forceinline__ void topkInsert(Key& mine, Key& kth, const Key x, int k, int lane) {
  const unsigned FULL = 0xffffffffu;
  const bool gt = keyLess(x, mine);  // monotone over the lanes: true from the insertion point on
  const unsigned g = __ballot_sync(FULL, gt);
  const Key up = {__shfl_up_sync(FULL, mine.d, 1), __shfl_up_sync(FULL, mine.i, 1)};
  if (gt) mine = (lane > 0 && ((g >> (lane - 1)) & 1u)) ? up : x;
  kth.d = __shfl_sync(FULL, mine.d, k - 1);
  kth.i = __shfl_sync(FULL, mine.i, k - 1);
}

}  // namespace hppb200
