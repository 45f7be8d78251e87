// Tensor-core candidate pass for the brute-force WeighedDistance k-NN (north_star: "k-NN as a tensor-core GEMM with
// an exact FP64 re-rank"; reference src/nearest-neighbor/basic.cc:84-126, src/weighed-distance.cc:272-289).
//
// For vector-space models the squared WeighedDistance is ||a'||^2 + ||b'||^2 - 2 a'.b' with a' = w o a: a GEMM of
// the query tile against the node stream.  It runs on the 5th-generation tensor cores (tcgen05.mma kind::tf32,
// M = 128 queries x N = 256 nodes, accumulators in TMEM) with a three-term TF32 split so that the approximate
// squared distance carries an absolute error far below the gaps that matter:
//   query row  [ ah | ah | al | na_hi na_lo 1 1 ]      node row  [ -2bh | -2bl | -2bh | 1 1 nb_hi nb_lo ]
// (h/l = TF32 head and tail of the FP64 value, na/nb the FP64 squared norms), i.e. K = 3 nq + 4 slots padded to 8 kc.
// The accumulator IS the approximate squared distance, so the epilogue is one compare per (query, node) against the
// query's threshold; survivors are appended to the query's candidate list.  The threshold is an exact upper bound of
// the k-th distance: the k-th exact distance over a PREFIX of the nodes (first 4096 by the exact scan, then the
// candidates of a 16x larger prefix, ...), inflated by the error bound eps = 2^-16 (na + max nb).  Every node whose
// exact distance is <= the k-th exact distance therefore survives, and the final result is the exact FP64 top-k of
// the survivors by the same (distance, node index) key and the same operation order as the exact scan: identical
// indices and distances, bit for bit.  A query whose list overflows is redone by the exact scan.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "dev_knn.cuh"

namespace hppb200 {

constexpr int TC_M = 128;            // queries per A tile (UMMA M)
constexpr int TC_N = 256;            // nodes per B tile (UMMA N)
constexpr int TC_CAP = 1024;         // candidates kept per query and stage
constexpr int TC_EPI_WARPS = 8;      // two groups of 4 warps, one per accumulator buffer
constexpr int TC_THREADS = (TC_EPI_WARPS + 2) * 32;  // + bulk-copy producer warp + MMA issuer warp
constexpr int TC_BSTAGES = 2;        // node-tile ring
constexpr int TC_MAX_KC = 6;
constexpr int TC_MAX_QT = 8;
constexpr size_t TC_PREFIX0 = 4096;  // nodes of the exact bootstrap scan
constexpr size_t TC_MIN_NODES = 16384;

// float index of (row r, slot s) inside a tile in the UMMA K-major no-swizzle canonical layout: 8 x 16-byte core
// matrices, the two 16-byte halves of a K = 8 step 128 B apart (LBO), 8-row groups kc * 256 B apart (SBO)
__host__ __device__ __forceinline__ size_t tcIndex(int r, int s, int kc) {
  return (size_t)(r >> 3) * (size_t)(kc * 64) + (size_t)(s >> 2) * 32 + (size_t)(r & 7) * 4 + (size_t)(s & 3);
}

__device__ __forceinline__ float toTf32(double x) {
  const float f = (float)x;
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(f));
  return __uint_as_float(u);
}

// one thread per operand row: FP64 configuration -> TF32 split operand row.  side 0 = query (A), 1 = node (B).
// rows [row0, rowsEnd) are written; rows >= nvalid are padding (norm slot 1e30: never below a threshold).
__global__ void knnTcPrepKernel(const double* __restrict__ src, const double* __restrict__ scale, int nq, int kc, int side,
                                int rowsPerTile, size_t row0, size_t rowsEnd, size_t nvalid, float* __restrict__ dst,
                                unsigned* __restrict__ maxNorm) {
  const size_t row = row0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= rowsEnd) return;
  float v[8 * TC_MAX_KC];
#pragma unroll
  for (int s = 0; s < 8 * TC_MAX_KC; ++s) v[s] = 0.f;
  if (row < nvalid) {
    double norm = 0;
    for (int c = 0; c < nq; ++c) {
      const double x = scale[c] * src[row * (size_t)nq + c];
      norm = fma(x, x, norm);
      const float h = toTf32(x);
      const float l = toTf32(x - (double)h);
      if (side == 0) { v[c] = h; v[nq + c] = h; v[2 * nq + c] = l; }
      else { v[c] = -2.f * h; v[nq + c] = -2.f * l; v[2 * nq + c] = -2.f * h; }
    }
    const float nh = toTf32(norm);
    const float nl = toTf32(norm - (double)nh);
    const int o = 3 * nq;
    if (side == 0) { v[o] = nh; v[o + 1] = nl; v[o + 2] = 1.f; v[o + 3] = 1.f; }
    else {
      v[o] = 1.f; v[o + 1] = 1.f; v[o + 2] = nh; v[o + 3] = nl;
      atomicMax(maxNorm, __float_as_uint(__double2float_ru(norm)));
    }
  } else {
    v[3 * nq + (side == 0 ? 0 : 2)] = toTf32(1e30);
  }
  const size_t tile = row / (size_t)rowsPerTile;
  const int r = (int)(row % (size_t)rowsPerTile);
  float* t = dst + tile * (size_t)rowsPerTile * (size_t)(8 * kc);
  for (int c = 0; c < 2 * kc; ++c)
    *(float4*)(t + tcIndex(r, 4 * c, kc)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
}

// ---- PTX helpers ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  const long long t0 = clock64();
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (clock64() - t0 > 4000000000ll) __trap();  // ~2 s: a protocol error must not hang the device
  }
}
__device__ __forceinline__ void mbarArrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbarExpectTx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulkCopy(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void tcFenceBefore() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcFenceAfter() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcCommit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tcMma(uint32_t tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 consecutive accumulator columns of this thread's TMEM lane (= query row), complete on return
__device__ __forceinline__ void tcLoad32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor: start >> 4 at [0,14), leading
// byte offset >> 4 at [16,30), stride byte offset >> 4 at [32,46), version 1 at [46,48), layout type 0 at [61,64))
__device__ __forceinline__ uint64_t tcSmemDesc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) |
         (1ull << 46);
}

struct TcFilterArgs {
  const float* btiles;   // node operand tiles
  const float* atiles;   // query operand tiles of this launch (nqt tiles)
  const float* thr;      // [nqt * 128] thresholds (padding rows: -1)
  int* cnt;              // [nqt * 128] candidates found per query (may exceed TC_CAP)
  int* cand;             // [nqt * 128][TC_CAP] node indices
  float* dump;           // debug: accumulator of (query tile 0, node tile 0), [128][256], or null
  unsigned ntiles;       // node tiles of the prefix
  int nqt, kc;
  int swapOffsets;       // debug: exchange the leading/stride byte offsets of the descriptors
};

// Persistent grid, one CTA per SM.  Warp 8 lane 0: bulk copies (query tiles once, node tiles through a ring);
// warp 9 lane 0: tcgen05.mma issue, kc K = 8 steps per (node tile, query tile) into one of two 256-column TMEM
// accumulators; warps 0-3 / 4-7: epilogue of accumulator 0 / 1 (thread = TMEM lane = query row).
__global__ void __launch_bounds__(TC_THREADS, 1) knnTcFilterKernel(TcFilterArgs A) {
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* bars = (uint64_t*)smem;  // 0: a_full, 1..2: b_full, 3..4: b_empty, 5..6: acc_full, 7..8: acc_empty
  uint32_t* tmemPtr = (uint32_t*)(smem + 128);
  const uint32_t aBytes = (uint32_t)(TC_M * A.kc * 32), bBytes = (uint32_t)(TC_N * A.kc * 32);
  unsigned char* sA = smem + 256;
  unsigned char* sB = sA + (size_t)A.nqt * aBytes;
  float* sThr = (float*)(sB + (size_t)TC_BSTAGES * bBytes);
  const uint32_t barBase = smemAddr(bars);
  auto bar = [&](int i) { return barBase + 8u * (uint32_t)i; };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbarInit(bar(0), 1);
    for (int s = 0; s < TC_BSTAGES; ++s) { mbarInit(bar(1 + s), 1); mbarInit(bar(3 + s), 1); }
    for (int b = 0; b < 2; ++b) { mbarInit(bar(5 + b), 1); mbarInit(bar(7 + b), 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  for (int x = threadIdx.x; x < A.nqt * TC_M; x += TC_THREADS) sThr[x] = A.thr[x];
  if (warp == TC_EPI_WARPS) {  // the whole tensor memory of the SM: two 128 x 256 FP32 accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(tmemPtr)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcFenceBefore();
  __syncthreads();
  tcFenceAfter();
  const uint32_t tmemBase = *tmemPtr;

  if (warp == TC_EPI_WARPS) {
    if (lane == 0) {
      mbarExpectTx(bar(0), (uint32_t)A.nqt * aBytes);
      for (int qt = 0; qt < A.nqt; ++qt)
        bulkCopy(smemAddr(sA + (size_t)qt * aBytes), A.atiles + (size_t)qt * (aBytes / 4), aBytes, bar(0));
      unsigned it = 0;
      for (unsigned t = blockIdx.x; t < A.ntiles; t += gridDim.x, ++it) {
        const unsigned s = it % TC_BSTAGES, ph = (it / TC_BSTAGES) & 1u;
        mbarWait(bar(3 + s), ph ^ 1u);
        mbarExpectTx(bar(1 + s), bBytes);
        bulkCopy(smemAddr(sB + (size_t)s * bBytes), A.btiles + (size_t)t * (bBytes / 4), bBytes, bar(1 + s));
      }
    }
  } else if (warp == TC_EPI_WARPS + 1) {
    if (lane == 0) {
      // instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10), both
      // K-major, N >> 3 at bit 17, M >> 4 at bit 24
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_N >> 3) << 17) | ((uint32_t)(TC_M >> 4) << 24);
      const uint32_t lbo = A.swapOffsets ? (uint32_t)A.kc * 256u : 128u, sbo = A.swapOffsets ? 128u : (uint32_t)A.kc * 256u;
      mbarWait(bar(0), 0);
      unsigned it = 0, j = 0;
      for (unsigned t = blockIdx.x; t < A.ntiles; t += gridDim.x, ++it) {
        const unsigned s = it % TC_BSTAGES, ph = (it / TC_BSTAGES) & 1u;
        mbarWait(bar(1 + s), ph);
        tcFenceAfter();
        const uint32_t bAddr = smemAddr(sB + (size_t)s * bBytes);
        for (int qt = 0; qt < A.nqt; ++qt, ++j) {
          const unsigned buf = j & 1u, aph = (j >> 1) & 1u;
          mbarWait(bar(7 + buf), aph ^ 1u);
          tcFenceAfter();
          const uint32_t aAddr = smemAddr(sA + (size_t)qt * aBytes);
          for (int c = 0; c < A.kc; ++c)
            tcMma(tmemBase + buf * (uint32_t)TC_N, tcSmemDesc(aAddr + (uint32_t)c * 256u, lbo, sbo),
                  tcSmemDesc(bAddr + (uint32_t)c * 256u, lbo, sbo), idesc, c > 0 ? 1u : 0u);
          tcCommit(bar(5 + buf));
        }
        tcCommit(bar(3 + s));
      }
    }
  } else {
    const unsigned group = (unsigned)warp >> 2;               // accumulator buffer this warp drains
    const int row = (warp & 3) * 32 + lane;                    // TMEM lane = query row of the tile
    const uint32_t laneAddr = tmemBase + ((uint32_t)((warp & 3) * 32) << 16) + group * (uint32_t)TC_N;
    unsigned j = 0, use = 0;
    for (unsigned t = blockIdx.x; t < A.ntiles; t += gridDim.x) {
      for (int qt = 0; qt < A.nqt; ++qt, ++j) {
        if ((j & 1u) != group) continue;
        mbarWait(bar(5 + group), use & 1u);
        ++use;
        tcFenceAfter();
        const float thr = sThr[qt * TC_M + row];
        const int q = qt * TC_M + row;
        for (int cb = 0; cb < TC_N / 32; ++cb) {
          uint32_t r[32];
          __syncwarp();  // tcgen05.ld is warp-collective: reconverge after the wait loop / the candidate branch
          tcLoad32(laneAddr + (uint32_t)cb * 32u, r);
          bool any = false;
#pragma unroll
          for (int i = 0; i < 32; ++i) any |= __uint_as_float(r[i]) <= thr;
          if (any) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (__uint_as_float(r[i]) <= thr) {
                const int slot = atomicAdd(&A.cnt[q], 1);
                if (slot < TC_CAP) A.cand[(size_t)q * TC_CAP + slot] = (int)(t * (unsigned)TC_N + (unsigned)(cb * 32 + i));
              }
          }
          if (A.dump != nullptr && t == 0 && qt == 0) {
#pragma unroll
            for (int i = 0; i < 32; ++i) A.dump[(size_t)row * TC_N + cb * 32 + i] = __uint_as_float(r[i]);
          }
        }
        tcFenceBefore();
        mbarArrive(bar(7 + group));
      }
    }
  }
  tcFenceBefore();
  __syncthreads();
  if (warp == TC_EPI_WARPS) {
    tcFenceAfter();
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(512u) : "memory");
  }
}

// thresholds from the exact k-th distance over a node prefix: thr = d_k^2 (1 + 1e-9) + eps, eps = 2^-16 (na + max nb)
// rounded up to FP32; rows beyond nqueries get -1 (their operand rows are padding)
__global__ void knnTcThresholdKernel(const double* __restrict__ queries, const double* __restrict__ scale, int nq, int nqueries,
                                     int rows, const double* __restrict__ kdist, int k, const unsigned* __restrict__ maxNorm,
                                     float* __restrict__ thr) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= rows) return;
  if (q >= nqueries) { thr[q] = -1.f; return; }
  double na = 0;
  for (int c = 0; c < nq; ++c) { const double x = scale[c] * queries[(size_t)q * nq + c]; na = fma(x, x, na); }
  const double eps = ldexp(na + (double)__uint_as_float(*maxNorm), -16) + 1e-30;
  const double d = kdist[(size_t)q * k + (k - 1)];
  thr[q] = __double2float_ru(d * d * (1.0 + 1e-9) + eps);  // +inf when the prefix held fewer than k nodes
}

// one warp per query: exact FP64 WeighedDistance of the candidates (operation order of the exact scan), k smallest
// by (distance, node index).  Not final: tighten the threshold.  Final: write the result; a query whose candidate
// list overflowed is flagged (idx[q][0] = -2, appended to `flagged`) and redone by the exact scan.
__global__ void __launch_bounds__(128)
knnTcRerankKernel(ModelView mv, const double* __restrict__ nodes, const double* __restrict__ queries,
                  const double* __restrict__ scale, int nqueries, int k, const int* __restrict__ cnt, const int* __restrict__ cand,
                  const unsigned* __restrict__ maxNorm, float* __restrict__ thr, int finalStage, long long* __restrict__ idx,
                  double* __restrict__ dist, unsigned* __restrict__ nflagged, int* __restrict__ flagged) {
  extern __shared__ __align__(16) unsigned char rsm[];
  const Tables T = makeTables(mv.blob, mv);
  double* sw2 = (double*)rsm;
  double* sqAll = sw2 + mv.njoints;
  for (int j = threadIdx.x; j < mv.njoints; j += blockDim.x) { const double w = T.weights[j]; sw2[j] = w * w; }
  const unsigned FULL = 0xffffffffu;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.x * (blockDim.x / 32) + warp;
  double* sq = sqAll + (size_t)warp * mv.nq;
  if (q < nqueries)
    for (int c = lane; c < mv.nq; c += 32) sq[c] = queries[(size_t)q * mv.nq + c];
  __syncthreads();
  if (q >= nqueries) return;
  const int total = cnt[q];
  if (total > TC_CAP) {
    if (finalStage && lane == 0) {
      idx[(size_t)q * k] = -2;
      flagged[atomicAdd(nflagged, 1u)] = q;
    }
    return;  // (not final: the previous threshold stays, it is still an upper bound)
  }
  Key acc = {CUDART_INF, LLONG_MAX}, ak = acc;
  for (int f0 = 0; f0 < total; f0 += 32) {
    const int f = f0 + lane;
    Key c = {CUDART_INF, LLONG_MAX};
    if (f < total) {
      const long long node = cand[(size_t)q * TC_CAP + f];
      double d1[1];
      weighedDistanceBatch<1, false>(T, mv, nodes + (size_t)node * mv.nq, sq, sw2, d1);
      c.d = d1[0]; c.i = node;
    }
    unsigned m = __ballot_sync(FULL, keyLess(c, ak));
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1u;
      const Key x = {__shfl_sync(FULL, c.d, src), __shfl_sync(FULL, c.i, src)};
      if (keyLess(x, ak)) topkInsert(acc, ak, x, k, lane);
    }
  }
  if (finalStage) {
    if (lane < k) {
      idx[(size_t)q * k + lane] = acc.i != LLONG_MAX ? acc.i : -1;
      dist[(size_t)q * k + lane] = acc.d;
    }
  } else if (lane == 0 && ak.i != LLONG_MAX) {
    double na = 0;
    for (int c = 0; c < mv.nq; ++c) { const double x = scale[c] * sq[c]; na = fma(x, x, na); }
    const double eps = ldexp(na + (double)__uint_as_float(*maxNorm), -16) + 1e-30;
    const float t = __double2float_ru(ak.d * ak.d * (1.0 + 1e-9) + eps);
    if (t < thr[q]) thr[q] = t;
  }
}

__global__ void knnTcGatherQueriesKernel(const double* __restrict__ queries, const int* __restrict__ list, int n, int nq,
                                         double* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x < n * nq) out[x] = queries[(size_t)list[x / nq] * nq + x % nq];
}
__global__ void knnTcScatterResultsKernel(const long long* __restrict__ si, const double* __restrict__ sd, const int* __restrict__ list,
                                          int n, int k, long long* __restrict__ idx, double* __restrict__ dist) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x < n * k) {
    const size_t o = (size_t)list[x / k] * k + x % k;
    idx[o] = si[x];
    dist[o] = sd[x];
  }
}

// ---- host side -------------------------------------------------------------------------------------------------
static size_t tcSmemBytes(int nqt, int kc) {
  return 256 + (size_t)nqt * TC_M * kc * 32 + (size_t)TC_BSTAGES * TC_N * kc * 32 + (size_t)nqt * TC_M * sizeof(float);
}
static int tcMaxQueryTiles(int kc) {
  int nqt = TC_MAX_QT;
  while (nqt > 0 && tcSmemBytes(nqt, kc) > 220 * 1024) --nqt;
  return nqt;
}

void knnTcInit(hppb200_knn* nn) {
  if (const char* v = std::getenv("HPPB200_KNN_TC")) nn->tc_enabled = std::atoi(v) != 0;
  if (const char* v = std::getenv("HPPB200_KNN_TC_MIN")) nn->tc_min_queries = std::max(1, std::atoi(v));
  nn->tc_kc = 0;
  if (nn->ff) return;
  const int nq = nn->model->nq;
  const int kc = (3 * nq + 4 + 7) / 8;
  if (kc > TC_MAX_KC || tcMaxQueryTiles(kc) < 1) return;
  nn->tc_kc = kc;
  // per configuration coordinate: the weight of its joint for rotations, 1 for translations (weighed-distance.cc:280-287)
  std::vector<double> scale((size_t)nq, 1.0);
  for (size_t j = 0; j < nn->model->joints.size(); ++j) {
    const DevJoint& J = nn->model->joints[j];
    if (J.type <= HPPB200_JOINT_REVOLUTE_Z) scale[(size_t)J.idx_q] = nn->model->weights[j];
  }
  nn->d_tc_scale.reserve((size_t)nq);
  HPPB200_CUDA_CHECK(cudaMemcpy(nn->d_tc_scale.p, scale.data(), (size_t)nq * sizeof(double), cudaMemcpyHostToDevice));
  nn->d_tc_misc.reserve(4);
  HPPB200_CUDA_CHECK(cudaMemset(nn->d_tc_misc.p, 0, 4 * sizeof(unsigned)));
}

void knnTcDestroy(hppb200_knn* nn) {
  if (nn->d_tc_nodes) cudaFree(nn->d_tc_nodes);
  nn->d_tc_nodes = nullptr;
}

bool knnTcEligible(const hppb200_knn* nn, size_t nqueries, int k, int32_t cc) {
  return nn->tc_enabled && nn->tc_kc > 0 && cc < 0 && k >= 1 && k <= 32 && nn->n >= TC_MIN_NODES &&
         nqueries >= (size_t)nn->tc_min_queries && nn->n < 0x7fffff00ull;
}

// convert the nodes appended since the last query into operand tiles (the last tile is padded up to its end)
static void tcSyncNodeOperand(hppb200_knn* nn) {
  const int kc = nn->tc_kc, nq = nn->model->nq;
  const size_t tileFloats = (size_t)TC_N * 8 * kc;
  const size_t needTiles = (nn->n + TC_N - 1) / TC_N;
  if (needTiles > nn->tc_cap_tiles) {
    const size_t ncap = std::max<size_t>((nn->cap + TC_N - 1) / TC_N, needTiles);
    float* nd = nullptr;
    HPPB200_CUDA_CHECK(cudaMalloc((void**)&nd, ncap * tileFloats * sizeof(float)));
    const size_t keep = nn->tc_n / TC_N;  // complete tiles already converted
    if (nn->d_tc_nodes && keep)
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(nd, nn->d_tc_nodes, keep * tileFloats * sizeof(float), cudaMemcpyDeviceToDevice, nn->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
    if (nn->d_tc_nodes) cudaFree(nn->d_tc_nodes);
    nn->d_tc_nodes = nd;
    nn->tc_cap_tiles = ncap;
    nn->tc_n = keep * TC_N;
  }
  if (nn->tc_n >= nn->n) return;
  const size_t row0 = (nn->tc_n / TC_N) * TC_N, rowsEnd = needTiles * TC_N;
  knnTcPrepKernel<<<(unsigned)((rowsEnd - row0 + 127) / 128), 128, 0, nn->stream>>>(nn->d_nodes, nn->d_tc_scale.p, nq, kc, 1, TC_N, row0,
                                                                                    rowsEnd, nn->n, nn->d_tc_nodes, nn->d_tc_misc.p);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  nn->tc_n = nn->n;
}

static void tcLaunchFilter(hppb200_knn* nn, size_t prefixNodes, int nqtTotal, float* dump, int swapOffsets) {
  const int kc = nn->tc_kc;
  const int perLaunch = tcMaxQueryTiles(kc);
  const unsigned ntiles = (unsigned)((prefixNodes + TC_N - 1) / TC_N);
  for (int qt0 = 0; qt0 < nqtTotal; qt0 += perLaunch) {
    TcFilterArgs a;
    a.nqt = std::min(perLaunch, nqtTotal - qt0);
    a.kc = kc;
    a.btiles = nn->d_tc_nodes;
    a.atiles = nn->d_tc_queries.p + (size_t)qt0 * TC_M * 8 * kc;
    a.thr = nn->d_tc_thr.p + (size_t)qt0 * TC_M;
    a.cnt = nn->d_tc_cnt.p + (size_t)qt0 * TC_M;
    a.cand = nn->d_tc_cand.p + (size_t)qt0 * TC_M * TC_CAP;
    a.dump = qt0 == 0 ? dump : nullptr;
    a.ntiles = ntiles;
    a.swapOffsets = swapOffsets;
    const size_t smemBytes = tcSmemBytes(a.nqt, kc);
    HPPB200_CUDA_CHECK(cudaFuncSetAttribute(knnTcFilterKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
    const unsigned grid = std::min<unsigned>((unsigned)nn->num_sms, ntiles);
    knnTcFilterKernel<<<grid, TC_THREADS, smemBytes, nn->stream>>>(a);
    HPPB200_CUDA_CHECK(cudaGetLastError());
  }
}

static void tcPrepQueries(hppb200_knn* nn, const double* d_q, size_t nqueries, int nqtTotal) {
  const int kc = nn->tc_kc, nq = nn->model->nq;
  const size_t rows = (size_t)nqtTotal * TC_M;
  nn->d_tc_queries.reserve(rows * 8 * kc);
  nn->d_tc_thr.reserve(rows);
  nn->d_tc_cnt.reserve(rows);
  nn->d_tc_cand.reserve(rows * TC_CAP);
  nn->d_tc_flagged.reserve(rows);
  knnTcPrepKernel<<<(unsigned)((rows + 127) / 128), 128, 0, nn->stream>>>(d_q, nn->d_tc_scale.p, nq, kc, 0, TC_M, 0, rows, nqueries,
                                                                          nn->d_tc_queries.p, nn->d_tc_misc.p);
  HPPB200_CUDA_CHECK(cudaGetLastError());
}

void knnTcRun(hppb200_knn* nn, const double* d_q, size_t nqueries, int k, long long* d_i, double* d_d) {
  const int nq = nn->model->nq;
  const int nqtTotal = (int)((nqueries + TC_M - 1) / TC_M);
  const size_t rows = (size_t)nqtTotal * TC_M;
  tcSyncNodeOperand(nn);
  tcPrepQueries(nn, d_q, nqueries, nqtTotal);
  // bootstrap: exact k-th distance over the first TC_PREFIX0 nodes
  nn->d_tc_kdist.reserve(nqueries * (size_t)k);
  nn->d_tc_kidx.reserve(nqueries * (size_t)k);
  knnExactRun(nn, d_q, nqueries, k, -1, TC_PREFIX0, nn->d_tc_kidx.p, nn->d_tc_kdist.p);
  knnTcThresholdKernel<<<(unsigned)((rows + 255) / 256), 256, 0, nn->stream>>>(d_q, nn->d_tc_scale.p, nq, (int)nqueries, (int)rows,
                                                                                nn->d_tc_kdist.p, k, nn->d_tc_misc.p, nn->d_tc_thr.p);
  // prefixes growing geometrically (<= 32x per stage) up to the whole roadmap
  const double ratio = (double)nn->n / (double)TC_PREFIX0;
  const int nstages = ratio <= 32.0 ? 1 : (ratio <= 1024.0 ? 2 : 3);
  const size_t rerankSmem = ((size_t)nn->model->njoints + 4 * (size_t)nq) * sizeof(double);
  HPPB200_CUDA_CHECK(cudaMemsetAsync(nn->d_tc_misc.p + 1, 0, sizeof(unsigned), nn->stream));
  for (int s = 1; s <= nstages; ++s) {
    size_t prefix = nn->n;
    if (s < nstages) {
      prefix = (size_t)((double)TC_PREFIX0 * std::pow(ratio, (double)s / nstages));
      prefix = std::min(nn->n, (prefix + TC_N - 1) / TC_N * TC_N);
    }
    HPPB200_CUDA_CHECK(cudaMemsetAsync(nn->d_tc_cnt.p, 0, rows * sizeof(int), nn->stream));
    tcLaunchFilter(nn, prefix, nqtTotal, nullptr, 0);
    knnTcRerankKernel<<<(unsigned)((nqueries + 3) / 4), 128, rerankSmem, nn->stream>>>(
        nn->view, nn->d_nodes, d_q, nn->d_tc_scale.p, (int)nqueries, k, nn->d_tc_cnt.p, nn->d_tc_cand.p, nn->d_tc_misc.p, nn->d_tc_thr.p,
        s == nstages ? 1 : 0, d_i, d_d, nn->d_tc_misc.p + 1, nn->d_tc_flagged.p);
    HPPB200_CUDA_CHECK(cudaGetLastError());
  }
  // queries whose final candidate list overflowed: exact scan
  unsigned nflagged = 0;
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&nflagged, nn->d_tc_misc.p + 1, sizeof(unsigned), cudaMemcpyDeviceToHost, nn->stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
  nn->tc_last_used = true;
  nn->tc_last_stages = nstages;
  nn->tc_last_overflow = nflagged;
  nn->tc_last_rows = rows;
  if (nflagged) {
    DeviceBuffer<double> gq, gd;
    DeviceBuffer<long long> gi;
    gq.reserve((size_t)nflagged * nq);
    gd.reserve((size_t)nflagged * k);
    gi.reserve((size_t)nflagged * k);
    knnTcGatherQueriesKernel<<<(nflagged * (unsigned)nq + 255) / 256, 256, 0, nn->stream>>>(d_q, nn->d_tc_flagged.p, (int)nflagged, nq, gq.p);
    knnExactRun(nn, gq.p, nflagged, k, -1, nn->n, gi.p, gd.p);
    knnTcScatterResultsKernel<<<(nflagged * (unsigned)k + 255) / 256, 256, 0, nn->stream>>>(gi.p, gd.p, nn->d_tc_flagged.p, (int)nflagged, k,
                                                                                           d_i, d_d);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
  }
}

void setLastError(const std::string& s);

}  // namespace hppb200

using namespace hppb200;

extern "C" {

int hppb200_knn_tc_stats(hppb200_knn* nn, int64_t* out) {
  try {
    if (!nn || !out) throw std::invalid_argument("null argument");
    HPPB200_CUDA_CHECK(cudaSetDevice(nn->device));
    out[0] = nn->tc_last_used ? 1 : 0;
    out[1] = nn->tc_last_stages;
    out[2] = (int64_t)nn->tc_last_overflow;
    out[3] = 0;
    if (nn->tc_last_used && nn->d_tc_cnt.p) {  // candidates of the last stage, summed on the host
      std::vector<int> h(nn->tc_last_rows);
      HPPB200_CUDA_CHECK(cudaMemcpy(h.data(), nn->d_tc_cnt.p, h.size() * sizeof(int), cudaMemcpyDeviceToHost));
      for (int c : h) out[3] += c > 0 ? c : 0;
    }
    return HPPB200_OK;
  } catch (const std::exception& e) {
    setLastError(e.what());
    return std::strncmp(e.what(), "CUDA", 4) == 0 ? HPPB200_ERR_CUDA : HPPB200_ERR_INVALID_ARGUMENT;
  }
}

int hppb200_knn_tc_debug_tile(hppb200_knn* nn, const double* queries, size_t nqueries, int swap_offsets, float* out) {
  try {
    if (!nn || !queries || !out || nqueries == 0 || nqueries > TC_M) throw std::invalid_argument("bad argument");
    if (nn->tc_kc == 0 || nn->n == 0) throw std::invalid_argument("tensor-core pass not available for this model / empty mirror");
    HPPB200_CUDA_CHECK(cudaSetDevice(nn->device));
    const size_t nq = (size_t)nn->model->nq;
    nn->d_queries.reserve(nqueries * nq);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(nn->d_queries.p, queries, nqueries * nq * sizeof(double), cudaMemcpyHostToDevice, nn->stream));
    tcSyncNodeOperand(nn);
    tcPrepQueries(nn, nn->d_queries.p, nqueries, 1);
    DeviceBuffer<float> dump;
    dump.reserve((size_t)TC_M * TC_N);
    HPPB200_CUDA_CHECK(cudaMemsetAsync(nn->d_tc_cnt.p, 0, TC_M * sizeof(int), nn->stream));
    HPPB200_CUDA_CHECK(cudaMemsetAsync(nn->d_tc_thr.p, 0xff, TC_M * sizeof(float), nn->stream));  // NaN: nothing passes
    tcLaunchFilter(nn, std::min<size_t>(nn->n, TC_N), 1, dump.p, swap_offsets);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(out, dump.p, (size_t)TC_M * TC_N * sizeof(float), cudaMemcpyDeviceToHost, nn->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
    return HPPB200_OK;
  } catch (const std::invalid_argument& e) {
    setLastError(e.what());
    return HPPB200_ERR_INVALID_ARGUMENT;
  } catch (const std::exception& e) {
    setLastError(e.what());
    return std::strncmp(e.what(), "CUDA", 4) == 0 ? HPPB200_ERR_CUDA : HPPB200_ERR_RUNTIME;
  }
}

}  // extern "C"
