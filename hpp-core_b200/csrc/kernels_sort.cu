// Longest-edge-first order for the wavefront engines.
//
// A batch of continuous validations ends when its LONGEST chain of dependent samples ends (Progressive on random
// UR5 edges: up to 540 samples, one per round).  Handing the edges out in input order lets a long edge start late
// (when a slot frees up around round 170) and stretch the batch to 880 rounds, most of them nearly empty.  The
// number of samples grows with the path length L, so the engines pull edges in order of decreasing L
// (longest-processing-time-first): the long chains start in round 0.  Which slot or round an edge runs in never
// reaches a result; outputs stay indexed by the caller's edge index.
#include <cub/device/device_radix_sort.cuh>

#include "engine.hpp"

namespace hppb200 {

__global__ void iotaKernel(uint32_t* v, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = (uint32_t)i;
}

// d_perm[k] = index of the edge with the k-th largest L.  L >= 0 (a distance), so the IEEE bit pattern orders
// like the value.  Scratch comes from ws->d_sort.  Asynchronous on ws->stream.
const uint32_t* sortEdgesByLength(Workspace* ws, const double* d_L, size_t n) {
  size_t tmpBytes = 0;
  const unsigned long long* keysIn = (const unsigned long long*)d_L;
  cub::DeviceRadixSort::SortPairsDescending(nullptr, tmpBytes, keysIn, (unsigned long long*)nullptr, (const uint32_t*)nullptr,
                                            (uint32_t*)nullptr, (int)n, 0, 64, ws->stream);
  const size_t oKeys = 0, oIdx = oKeys + ((n * 8 + 255) & ~size_t(255)), oPerm = oIdx + ((n * 4 + 255) & ~size_t(255));
  const size_t oTmp = oPerm + ((n * 4 + 255) & ~size_t(255));
  ws->d_sort.reserve(oTmp + tmpBytes + 256);
  unsigned char* base = ws->d_sort.p;
  uint32_t* idx = (uint32_t*)(base + oIdx);
  uint32_t* perm = (uint32_t*)(base + oPerm);
  iotaKernel<<<(unsigned)((n + 255) / 256), 256, 0, ws->stream>>>(idx, n);
  HPPB200_CUDA_CHECK(cub::DeviceRadixSort::SortPairsDescending(base + oTmp, tmpBytes, keysIn, (unsigned long long*)(base + oKeys), idx,
                                                               perm, (int)n, 0, 64, ws->stream));
  ws->stats.kernel_launches += 2;
  return perm;
}

}  // namespace hppb200
