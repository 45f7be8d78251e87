// Batched configuration validation: JointBoundValidation + CollisionValidation for many
// configurations at once (reference src/config-validations.cc:40-49,
// src/joint-bound-validation.cc:45-80, src/collision-validation.cc:48-95,
// src/obstacle-user.cc:42-51).
//
// One CTA validates CFG_PER_BLOCK configurations:
//   1. the model blob (joint/geometry/pair tables) is staged global->shared with one TMA
//      bulk copy;
//   2. FK phase: one thread per configuration walks the joint chain in registers and
//      writes the geometry placements oMg into shared memory, SoA over configurations
//      (lane = configuration => conflict-free);
//   3. pair phase: warps pull (pair, 32-configuration group) work items from a shared
//      counter in increasing pair order.  All lanes of a warp run the SAME pair (uniform
//      shape-kind dispatch), on 32 different configurations.  The first colliding pair of
//      the reference's sequential loop is the MINIMUM colliding index: atomicMin into
//      shared memory; a lane skips a pair whose index is already above its configuration's
//      current minimum (the reference's early exit).
// FP64 throughout; bound by the FP64 pipe, not HBM (53 algorithmic bytes per configuration).
#include "dev_model.cuh"

namespace hppb200 {

constexpr int CFG_PER_BLOCK = 64;
constexpr int CFG_THREADS = 256;
constexpr int MODE_VALIDATE = 0, MODE_DISTANCES = 1, MODE_PLACEMENTS = 2;

struct ConfigSmemLayout {
  uint32_t off_blob, off_store, off_first, off_bound, total;
};

static ConfigSmemLayout configSmemLayout(const ModelView& mv, bool tree) {
  ConfigSmemLayout L;
  uint32_t o = 16;  // [0,8) mbarrier, [8,12) work counter
  L.off_blob = o;
  o += mv.blob_bytes;
  L.off_store = o;
  uint32_t slots = (uint32_t)mv.ngeoms + (tree ? (uint32_t)mv.njoints : 0u);
  o += slots * 12u * CFG_PER_BLOCK * (uint32_t)sizeof(double);
  L.off_first = o;
  o += CFG_PER_BLOCK * sizeof(int);
  L.off_bound = o;
  o += CFG_PER_BLOCK * sizeof(int);
  L.total = o;
  return L;
}

template <bool TREE>
__global__ void __launch_bounds__(CFG_THREADS)
validateConfigsKernel(ModelView mv, ConfigSmemLayout L, const double* __restrict__ q, size_t n, uint32_t flags, int mode,
                      uint8_t* __restrict__ valid, int32_t* __restrict__ first, uint8_t* __restrict__ reason,
                      double* __restrict__ aux, DeviceStats* stats) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  int* next = (int*)(smem + 8);
  unsigned char* sblob = smem + L.off_blob;
  double* store = (double*)(smem + L.off_store);
  int* firstHit = (int*)(smem + L.off_first);
  int* boundHit = (int*)(smem + L.off_bound);

  stageBlobTMA(sblob, mv.blob, mv.blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);
  const size_t base = (size_t)blockIdx.x * CFG_PER_BLOCK;
  const int tid = threadIdx.x;
  if (tid == 0) *next = 0;

  // ---- phase 1: joint bounds + FK, one thread per configuration -------------------
  if (tid < CFG_PER_BLOCK) {
    firstHit[tid] = INT_MAX;
    boundHit[tid] = -1;
    const size_t i = base + tid;
    if (i < n) {
      const double* qc = q + i * (size_t)mv.nq;
      if (flags & HPPB200_CHECK_JOINT_BOUNDS) {
        for (int k = 0; k < mv.nq; ++k) {
          const double lo = T.lower[k], hi = T.upper[k];
          if (lo > hi) continue;  // unbounded rank (e.g. quaternion)
          if (qc[k] < lo || hi < qc[k]) { boundHit[tid] = k; break; }
        }
      }
      if (boundHit[tid] < 0) forwardKinematics<TREE>(T, mv, [qc](int k) { return qc[k]; }, store + tid, CFG_PER_BLOCK);
    }
  }
  __syncthreads();

  if (mode == MODE_PLACEMENTS) {
    // aux: [n][ngeoms][12]
    for (int e = tid; e < CFG_PER_BLOCK * mv.ngeoms * 12; e += CFG_THREADS) {
      const int c = e / (mv.ngeoms * 12), r = e % (mv.ngeoms * 12);
      if (base + c < n) aux[(base + c) * (size_t)(mv.ngeoms * 12) + r] = store[(size_t)r * CFG_PER_BLOCK + c];
    }
    return;
  }

  // ---- phase 2: narrow phase, warps pull (pair, 32-config group) items --------------
  const int lane = tid & 31;
  constexpr int GROUPS = CFG_PER_BLOCK / 32;
  const int nitems = mv.ncpairs * GROUPS;
  GjkCounters cnt = {0u, 0u, 0u};
  unsigned tests = 0;
  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(next, 1);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= nitems) break;
    const int p = item / GROUPS;
    const int c = (item % GROUPS) * 32 + lane;
    const size_t i = base + c;
    if (i >= n || boundHit[c] >= 0) continue;
    if (mode == MODE_VALIDATE && *(volatile int*)&firstHit[c] < p) continue;  // reference early exit
    const DevPair pr = T.cpairs[p];
    const double d = pairLowerBound(T, pr, store + c, CFG_PER_BLOCK, mv.vertices, cnt);
    ++tests;
    if (mode == MODE_DISTANCES) aux[i * (size_t)mv.ncpairs + p] = d;
    else if (d <= HPPB200_COLLISION_THRESHOLD) atomicMin(&firstHit[c], p);
  }
  __syncthreads();

  if (mode == MODE_VALIDATE && tid < CFG_PER_BLOCK) {
    const size_t i = base + tid;
    if (i < n) {
      const int b = boundHit[tid], h = firstHit[tid];
      const int rs = b >= 0 ? HPPB200_REASON_JOINT_BOUND : (h != INT_MAX ? HPPB200_REASON_COLLISION : HPPB200_REASON_VALID);
      valid[i] = rs == HPPB200_REASON_VALID ? 1 : 0;
      if (first) first[i] = b >= 0 ? b : (h != INT_MAX ? h : -1);
      if (reason) reason[i] = (uint8_t)rs;
    }
  }
  // statistics: warp-reduce then one atomic per warp
  unsigned t = __reduce_add_sync(0xffffffffu, tests);
  unsigned gi = __reduce_add_sync(0xffffffffu, cnt.iterations);
  unsigned gc = __reduce_add_sync(0xffffffffu, cnt.cap_hits);
  if (lane == 0 && stats) {
    if (t) atomicAdd(&stats->pair_tests, (unsigned long long)t);
    if (gi) atomicAdd(&stats->gjk_iterations, (unsigned long long)gi);
    if (gc) atomicAdd(&stats->gjk_cap_hits, (unsigned long long)gc);
  }
}

static void launchConfigKernel(Workspace* ws, const double* d_q, size_t n, uint32_t flags, int mode, uint8_t* d_valid,
                               int32_t* d_first, uint8_t* d_reason, double* d_aux) {
  if (n == 0) return;
  const bool tree = modelIsTree(ws->model);
  ConfigSmemLayout L = configSmemLayout(ws->view, tree);
  if (L.total > 227u * 1024u)
    throw std::runtime_error("model too large for the shared-memory configuration kernel (" + std::to_string(L.total) + " B)");
  const unsigned grid = (unsigned)((n + CFG_PER_BLOCK - 1) / CFG_PER_BLOCK);
  auto kern = tree ? validateConfigsKernel<true> : validateConfigsKernel<false>;
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total));
  kern<<<grid, CFG_THREADS, L.total, ws->stream>>>(ws->view, L, d_q, n, flags, mode, d_valid, d_first, d_reason, d_aux,
                                                  ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  if (mode == MODE_VALIDATE) ws->stats.configs += n;
}

void launchValidateConfigs(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid,
                           int32_t* d_first, uint8_t* d_reason) {
  launchConfigKernel(ws, d_q, n, flags, MODE_VALIDATE, d_valid, d_first, d_reason, nullptr);
}
void launchGeometryPlacements(Workspace* ws, const double* d_q, size_t n, double* d_oMg) {
  launchConfigKernel(ws, d_q, n, 0, MODE_PLACEMENTS, nullptr, nullptr, nullptr, d_oMg);
}
void launchPairDistances(Workspace* ws, const double* d_q, size_t n, double* d_dist) {
  launchConfigKernel(ws, d_q, n, 0, MODE_DISTANCES, nullptr, nullptr, nullptr, d_dist);
}


// FP64 FMA peak: 8 independent DFMA chains per thread, 4096 steps, enough CTAs to fill the chip.
__global__ void __launch_bounds__(256) fp64PeakKernel(double* out, double a, double b, int iters) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; ++i) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

double measureFp64Peak(int device) {
  HPPB200_CUDA_CHECK(cudaSetDevice(device));
  cudaDeviceProp prop;
  HPPB200_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 14;
  double* out = nullptr;
  HPPB200_CUDA_CHECK(cudaMalloc((void**)&out, (size_t)blocks * threads * sizeof(double)));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int rep = 0; rep < 6; ++rep) {
    cudaEventRecord(e0);
    fp64PeakKernel<<<blocks, threads>>>(out, 0.999999, 1e-7, iters);
    cudaEventRecord(e1);
    HPPB200_CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * 8 * (double)iters * blocks * threads;
    if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  return best;
}

}  // namespace hppb200
