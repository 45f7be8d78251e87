// Tensor-core candidate pass for the brute-force WeighedDistance k-NN (north_star: "k-NN as a tensor-core GEMM with
// an exact FP64 re-rank"; reference src/nearest-neighbor/basic.cc:84-126, src/weighed-distance.cc:272-289).
//
// For vector-space models the squared WeighedDistance is ||a'||^2 + ||b'||^2 - 2 a'.b' with a' = w o a: a GEMM of
// the query tile against the node stream.  It runs on the 5th-generation tensor cores (tcgen05.mma kind::tf32,
// M = 128 queries x N = 128 nodes, four accumulators in TMEM) with a three-term TF32 split so that the approximate
// squared distance carries an absolute error far below the gaps that matter:
//   query row  [ ah | ah | al | na_hi na_lo 1 1 -thr_hi -thr_lo ]   node row  [ -2bh | -2bl | -2bh | 1 1 nb_hi nb_lo 1 1 ]
// i.e. K = 3 nq + 6 slots padded to 8 kc (UR5: exactly 24).
// (h/l = TF32 head and tail of the FP64 value, na/nb the FP64 squared norms).
// Two more slots carry the query's threshold (-thr_hi, -thr_lo against 1, 1), so the accumulator IS the approximate
// squared distance minus the threshold and the epilogue is a sign test: 32 FP32 columns of a TMEM lane are OR-ed
// together; when the sign bit is set the GROUPS of 8 nodes holding a hit go to a list private to (query, CTA) (a
// shared-memory counter, no global atomics) and the re-rank looks at all 8 exactly.  The threshold is an exact upper bound of the k-th distance: the k-th
// exact distance over a PREFIX of the nodes (the first 512 by the re-rank kernel itself, then the candidates of a
// <= 16x larger prefix, ...), inflated by the error bound eps = 2^-15 (na + max nb) (2^-14 for models with more than 32 K slots).  Every node whose exact distance
// is <= the k-th exact distance therefore survives, and the final result is the exact FP64 top-k of the survivors by
// the same (distance, node index) key and the same operation order as the exact scan: identical indices and
// distances, bit for bit.  A query one of whose lists overflows is redone by the exact scan.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "dev_knn.cuh"

namespace hppb200 {

constexpr int TC_M = 128;            // queries per A tile (UMMA M)
constexpr int TC_N = 128;            // nodes per B tile (UMMA N)
constexpr int TC_NBUF = 4;           // TMEM accumulators (TC_NBUF * TC_N = 512 columns: the whole tensor memory)
constexpr int TC_CAPC = 16;          // candidates kept per (query, CTA) and stage when all SMs work (more per CTA otherwise)
constexpr int TC_EPI_WARPS = 4 * TC_NBUF;  // one group of 4 warps per accumulator buffer
constexpr int TC_ISSUERS = 2;        // warps issuing tcgen05.mma (lane 0 each), tile pairs dealt round-robin: ONE thread's serial
                                     // instruction stream was the pacing step of the kernel (profiles/r2b_knn_filter_ablation.txt)
constexpr int TC_THREADS = (TC_EPI_WARPS + 1 + TC_ISSUERS) * 32;  // + bulk-copy producer warp + MMA issuer warps
constexpr int TC_BSTAGES = 8;        // node-tile ring (at most; fewer when shared memory is short)
constexpr int TC_MAX_KC = 6;
constexpr int TC_MAX_QT = 8;
constexpr size_t TC_PREFIX0 = 512;   // nodes of the exact bootstrap (re-rank kernel over the first nodes)
constexpr double TC_RATIO = 16.0;    // prefix growth per stage (candidates per query and stage ~ k * ratio)
constexpr int TC_LIST = 2048;         // candidate groups a query may carry into the re-rank
constexpr int TC_KEYS = 2048;         // exact keys a query may keep for the sort path (k > 32)
constexpr int TC_MAX_K = 256;
constexpr float TC_THR_MAX = 1e29f;  // thresholds are clamped below the padding norm 1e30
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
// thr (queries only): per-row threshold folded into the operand, or null for 0.
__global__ void knnTcPrepKernel(const double* __restrict__ src, const double* __restrict__ scale, int nq, int kc, int side,
                                int rowsPerTile, size_t row0, size_t rowsEnd, size_t nvalid, float* __restrict__ dst,
                                unsigned* __restrict__ maxNorm, const float* __restrict__ thr) {
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
    if (side == 0) {
      v[o] = nh; v[o + 1] = nl; v[o + 2] = 1.f; v[o + 3] = 1.f;
      const float t = thr ? thr[row] : 0.f;  // accumulator = approximate squared distance - threshold
      const float th = toTf32((double)t);
      v[o + 4] = -th; v[o + 5] = -toTf32((double)t - (double)th);
    } else {
      v[o] = 1.f; v[o + 1] = 1.f; v[o + 2] = nh; v[o + 3] = nl; v[o + 4] = 1.f; v[o + 5] = 1.f;
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
  uint32_t done = 0, spins = 0;
  long long t0 = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if ((++spins & 0x3ffu) == 0) {  // a protocol error must not hang the device: trap after ~2 s
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > 4000000000ll) __trap();
    }
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
// one lane of the (converged) warp, the same one every time
__device__ __forceinline__ bool electOne() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
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
// the same with the descriptors given as their low words (start address and leading byte offset fields; advancing the
// start address by x bytes adds x >> 4) and the common high word: two moves instead of 64-bit shifts in the issuing thread
__device__ __forceinline__ void tcMmaLo(uint32_t tmem, uint32_t aLo, uint32_t bLo, uint32_t hi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %4, p;\n\t}" ::"r"(tmem),
      "r"(aLo), "r"(bLo), "r"(hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
template <int KC>
__device__ __forceinline__ void tcIssuePair(uint32_t tmem, uint32_t aLo, uint32_t bLo, uint32_t hi, uint32_t idesc) {
#pragma unroll
  for (int c = 0; c < KC; ++c) tcMmaLo(tmem, aLo + 16u * (uint32_t)c, bLo + 16u * (uint32_t)c, hi, idesc, c > 0 ? 1u : 0u);  // 256 B per K = 8 slot
}
// 32 consecutive accumulator columns of this thread's TMEM lane (= query row); valid after tcWaitLoad on the same array
__device__ __forceinline__ void tcLoad32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// all tcgen05.ld of this thread have landed; the registers are tied to the wait so that no use moves above it
__device__ __forceinline__ void tcWaitLoad(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]),
                 "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]),
                 "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]),
                 "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
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
  const float* atiles;   // query operand tiles of this launch (nqt tiles), thresholds folded in
  int* cnt;              // [nqt * 128][gstride] candidate groups found per (query, CTA) (may exceed capc)
  int* cand;             // [nqt * 128][gstride * TC_CAPC] candidate groups (8 consecutive nodes each), capc per CTA
  float* dump;           // debug: accumulator of (query tile 0, node tile 0), [128][256], or null
  unsigned ntiles;       // node tiles of the prefix
  int nqt, kc, bstages, gstride, capc;  // capc: list capacity per (query, CTA) of this launch
  int swapOffsets;       // debug: exchange the leading/stride byte offsets of the descriptors
  long long* prof;       // timing experiments: [grid][8] cycles (MMA thread: wait B, wait accumulator, issue, total; epilogue warp 0:
                         // wait accumulator, drain, total) or null
  int ablate;            // timing experiments only (results are wrong): 1 tcgen05.ld without the sign test, 2 no tcgen05.ld either,
                         // 3 = 2 + no tcgen05.mma (barrier protocol alone), 4 = 2 + one tcgen05.mma per tile pair
};

// Persistent grid, one CTA per SM.  Warp TC_EPI_WARPS lane 0: bulk copies (query tiles once, node tiles through a ring);
// next TC_ISSUERS warps, lane 0 each: tcgen05.mma issue, kc K = 8 steps per (node tile, query tile) into one of TC_NBUF
// 128-column TMEM accumulators; epilogue warps 4 g .. 4 g + 3 drain accumulator g (thread = TMEM lane = query row), two tcgen05.ld in
// flight per warp.
template <bool PROF>
__global__ void __launch_bounds__(TC_THREADS, 1) knnTcFilterKernel(TcFilterArgs A) {
  extern __shared__ __align__(128) unsigned char smem[];
  // barriers: 0 a_full, B_FULL.. b_full, B_EMPTY.. b_empty, ACC_FULL.. acc_full, ACC_EMPTY.. acc_empty
  constexpr int B_FULL = 1, B_EMPTY = B_FULL + TC_BSTAGES, ACC_FULL = B_EMPTY + TC_BSTAGES, ACC_EMPTY = ACC_FULL + TC_NBUF;
  static_assert((ACC_EMPTY + TC_NBUF) * 8 <= 384, "barrier block");
  uint64_t* bars = (uint64_t*)smem;
  uint32_t* tmemPtr = (uint32_t*)(smem + 384);
  const uint32_t aBytes = (uint32_t)(TC_M * A.kc * 32), bBytes = (uint32_t)(TC_N * A.kc * 32);
  unsigned char* sA = smem + 512;
  unsigned char* sB = sA + (size_t)A.nqt * aBytes;
  int* sCnt = (int*)(sB + (size_t)A.bstages * bBytes);  // candidates of this CTA per query row
  const uint32_t barBase = smemAddr(bars);
  auto bar = [&](int i) { return barBase + 8u * (uint32_t)i; };
  // (shuffles from lane 0: the compiler then keeps the warp index, and below the TMEM base, in uniform registers, which is
  // where tcgen05.mma / commit / ld take their operands from; per-thread values cost an R2UR waterfall loop per instruction)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const unsigned S = (unsigned)A.bstages;

  if (threadIdx.x == 0) {
    mbarInit(bar(0), 1);
    for (int s = 0; s < TC_BSTAGES; ++s) { mbarInit(bar(B_FULL + s), 1); mbarInit(bar(B_EMPTY + s), TC_ISSUERS); }
    for (int b = 0; b < TC_NBUF; ++b) { mbarInit(bar(ACC_FULL + b), 1); mbarInit(bar(ACC_EMPTY + b), 4); }  // one arrival per draining warp
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  for (int x = threadIdx.x; x < A.nqt * TC_M; x += TC_THREADS) sCnt[x] = 0;
  if (warp == TC_EPI_WARPS) {  // the whole tensor memory of the SM: TC_NBUF accumulators of 128 x TC_N FP32
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(tmemPtr)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcFenceBefore();
  __syncthreads();
  tcFenceAfter();
  const uint32_t tmemBase = __shfl_sync(0xffffffffu, *tmemPtr, 0);

  if (warp == TC_EPI_WARPS) {
    if (lane == 0) {
      mbarExpectTx(bar(0), (uint32_t)A.nqt * aBytes);
      for (int qt = 0; qt < A.nqt; ++qt)
        bulkCopy(smemAddr(sA + (size_t)qt * aBytes), A.atiles + (size_t)qt * (aBytes / 4), aBytes, bar(0));
      unsigned it = 0;
      for (unsigned t = blockIdx.x; t < A.ntiles; t += gridDim.x, ++it) {
        const unsigned s = it % S, ph = (it / S) & 1u;
        mbarWait(bar(B_EMPTY + s), ph ^ 1u);
        mbarExpectTx(bar(B_FULL + s), bBytes);
        bulkCopy(smemAddr(sB + (size_t)s * bBytes), A.btiles + (size_t)t * (bBytes / 4), bBytes, bar(B_FULL + s));
      }
    }
  } else if (warp > TC_EPI_WARPS) {
    // Issuing warp `me` of TC_ISSUERS takes the tile pairs j = me (mod TC_ISSUERS).  The WHOLE warp runs the loop on
    // warp-uniform values and one elected lane issues: the loop body is then a few dozen uniform-datapath instructions
    // (one thread's serial latency through this loop is what paces the tensor pipe at K = 8 kc).
    const unsigned me = (unsigned)(warp - (TC_EPI_WARPS + 1));
    // instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10), both
    // K-major, N >> 3 at bit 17, M >> 4 at bit 24
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_N >> 3) << 17) | ((uint32_t)(TC_M >> 4) << 24);
    const uint32_t lbo = A.swapOffsets ? (uint32_t)A.kc * 256u : 128u, sbo = A.swapOffsets ? 128u : (uint32_t)A.kc * 256u;
    const uint64_t d0 = tcSmemDesc(0, lbo, sbo);
    const uint32_t descHi = (uint32_t)(d0 >> 32), lboField = (uint32_t)d0;
    const uint32_t aLo0 = ((smemAddr(sA) & 0x3FFFFu) >> 4) | lboField, bLo0 = ((smemAddr(sB) & 0x3FFFFu) >> 4) | lboField;
    const uint32_t aStep = aBytes >> 4, bStep = bBytes >> 4;
    const int kcUse = A.ablate == 3 ? 0 : A.ablate == 4 ? 1 : A.kc;  // (timing experiments)
    mbarWait(bar(0), 0);
    unsigned it = 0, j = 0;
    long long pB = 0, pAcc = 0, pT0 = PROF ? clock64() : 0, c0 = 0;
    for (unsigned t = blockIdx.x; t < A.ntiles; t += gridDim.x, ++it) {
      const unsigned s = it % S, ph = (it / S) & 1u;
      if (PROF) c0 = clock64();
      mbarWait(bar(B_FULL + s), ph);
      if (PROF) pB += clock64() - c0;
      tcFenceAfter();
      const uint32_t bLo = bLo0 + s * bStep;
      uint32_t aLo = aLo0;
      for (int qt = 0; qt < A.nqt; ++qt, ++j, aLo += aStep) {
        if (j % TC_ISSUERS != me) continue;
        const unsigned buf = j % TC_NBUF;
        if (PROF) c0 = clock64();
        mbarWait(bar(ACC_EMPTY + buf), ((j / TC_NBUF) & 1u) ^ 1u);
        if (PROF) pAcc += clock64() - c0;
        tcFenceAfter();
        const uint32_t acc = tmemBase + buf * (uint32_t)TC_N;
        if (electOne()) {
          switch (kcUse) {
            case 1: tcIssuePair<1>(acc, aLo, bLo, descHi, idesc); break;
            case 2: tcIssuePair<2>(acc, aLo, bLo, descHi, idesc); break;
            case 3: tcIssuePair<3>(acc, aLo, bLo, descHi, idesc); break;
            case 4: tcIssuePair<4>(acc, aLo, bLo, descHi, idesc); break;
            case 5: tcIssuePair<5>(acc, aLo, bLo, descHi, idesc); break;
            case 6: tcIssuePair<6>(acc, aLo, bLo, descHi, idesc); break;
            default: break;
          }
          tcCommit(bar(ACC_FULL + buf));
        }
        __syncwarp();
      }
      if (electOne()) tcCommit(bar(B_EMPTY + s));  // (TC_ISSUERS arrivals free the stage: a commit covers the committing thread's MMAs)
      __syncwarp();
    }
    if (PROF && me == 0 && lane == 0) {
      const long long tot = clock64() - pT0;
      long long* P = A.prof + 8 * blockIdx.x;
      P[0] = pB; P[1] = pAcc; P[2] = tot - pB - pAcc; P[3] = tot;
    }
  } else {
    const unsigned group = (unsigned)warp >> 2;               // accumulator buffer this warp drains
    const int row = (warp & 3) * 32 + lane;                    // TMEM lane = query row of the tile
    const uint32_t laneAddr = tmemBase + ((uint32_t)((warp & 3) * 32) << 16) + group * (uint32_t)TC_N;
    unsigned use = 0;
    const bool profMe = PROF && threadIdx.x == 0;
    long long pW = 0, pT0 = profMe ? clock64() : 0, c0 = 0;
    // this group's tile pairs only: j = group, group + TC_NBUF, ... in the issuers' order (t outer, qt inner); stepping
    // through all pairs and skipping three of four cost the draining warps ~15 % of their issue slots
    unsigned t = blockIdx.x;
    int qt = (int)group;
    for (;; qt += TC_NBUF) {
      while (qt >= A.nqt) { qt -= A.nqt; t += gridDim.x; }
      if (t >= A.ntiles) break;
      {
        if (profMe) c0 = clock64();
        mbarWait(bar(ACC_FULL + group), use & 1u);
        if (profMe) pW += clock64() - c0;
        ++use;
        tcFenceAfter();
        const int q = qt * TC_M + row;
        // 32 columns: OR of the sign bits first; a set sign bit = approximate squared distance below the threshold
        auto process = [&](const uint32_t (&r)[32], int cb) {
          // (reduction trees of depth 3 + 1, not one chain of 32 dependent ORs: the chain's latency, 4 blocks per
          // accumulator, was what the draining warps spent their time on)
          uint32_t og[4];
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t t1 = r[8 * g] | r[8 * g + 1] | r[8 * g + 2], t2 = r[8 * g + 3] | r[8 * g + 4] | r[8 * g + 5];
            og[g] = (t1 | r[8 * g + 6]) | (t2 | r[8 * g + 7]);
          }
          const uint32_t any = (og[0] | og[1]) | (og[2] | og[3]);
          if ((int)any < 0) {  // rare: the groups of 8 nodes holding a hit go to the query's list of this CTA
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const uint32_t o = og[g];
              if ((int)o < 0) {
                const int slot = atomicAdd(&sCnt[q], 1);
                if (slot < A.capc)
                  A.cand[(size_t)q * A.gstride * TC_CAPC + (size_t)blockIdx.x * A.capc + slot] =
                      (int)(t * (unsigned)(TC_N / 8) + (unsigned)(cb * 4 + g));
              }
            }
          }
          if (A.dump != nullptr && t < 2 && qt == 0) {
#pragma unroll
            for (int i = 0; i < 32; ++i) A.dump[(size_t)row * (2 * TC_N) + t * TC_N + cb * 32 + i] = __uint_as_float(r[i]);
          }
          __syncwarp();  // tcgen05.ld / wait are warp-collective: reconverge after the candidate branch
        };
        uint32_t ra[32], rb[32];
        __syncwarp();
        if (A.ablate) {
          if (A.ablate == 1) {
            uint32_t acc = 0;
#pragma unroll
            for (int cb = 0; cb < TC_N / 32; cb += 2) {
              tcLoad32(laneAddr + (uint32_t)cb * 32u, ra);
              tcLoad32(laneAddr + (uint32_t)(cb + 1) * 32u, rb);
              tcWaitLoad(ra);
              tcWaitLoad(rb);
              acc |= ra[0] | rb[31];
            }
            if (acc == 0x12345u) sCnt[q] = 1;
          }
          tcFenceBefore();
          __syncwarp();
          if (lane == 0) mbarArrive(bar(ACC_EMPTY + group));
          continue;
        }
        // two loads in flight; the next pair is issued as soon as its registers are free, so that its latency runs
        // under the sign tests of the current pair (tcgen05.wait::ld waits for ALL loads of the thread: one wait per pair)
        tcLoad32(laneAddr, ra);
        tcLoad32(laneAddr + 32u, rb);
#pragma unroll
        for (int cb = 0; cb < TC_N / 32; cb += 2) {
          tcWaitLoad(ra);
          tcWaitLoad(rb);
          process(ra, cb);
          if (cb + 2 < TC_N / 32) tcLoad32(laneAddr + (uint32_t)(cb + 2) * 32u, ra);
          process(rb, cb + 1);
          if (cb + 3 < TC_N / 32) tcLoad32(laneAddr + (uint32_t)(cb + 3) * 32u, rb);
        }
        // every lane's loads have landed (tcgen05.wait::ld above); one arrival per warp: 128 arrivals on one barrier word
        // serialise in the shared-memory atomic unit and were the pacing step of the kernel (profiles/r2b_knn_filter_ablation.txt)
        tcFenceBefore();
        __syncwarp();
        if (lane == 0) mbarArrive(bar(ACC_EMPTY + group));
      }
    }
    if (profMe) {
      const long long tot = clock64() - pT0;
      long long* P = A.prof + 8 * blockIdx.x;
      P[4] = pW; P[5] = tot - pW; P[6] = tot; P[7] = (long long)use;
    }
  }
  tcFenceBefore();
  __syncthreads();
  for (int x = threadIdx.x; x < A.nqt * TC_M; x += TC_THREADS) A.cnt[(size_t)x * A.gstride + blockIdx.x] = sCnt[x];
  if (warp == TC_EPI_WARPS) {
    tcFenceAfter();
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(512u) : "memory");
  }
}

// one CTA per query: exact FP64 WeighedDistance of the candidates (operation order of the exact scan), k smallest
// by (distance, node index).  A candidate is a group of 8 consecutive nodes: the first `prefixNodes` nodes
// (bootstrap, cnt == null) or the query's per-CTA lists of the last filter launch, compacted into shared memory by
// warp 0; every warp then takes 4 groups per trip (lane = node), keeps its k best sorted across the lanes, and warp
// 0 merges.  Not final: the new threshold thr = d_k^2 (1 + 1e-6) + eps, eps = 2^-15 .. 2^-14 (na + max nb), rounded up, if
// it is tighter.  Final: write the result; a query with an overflowed list is flagged (idx[q][0] = -2, appended to
// `flagged`) and redone by the exact scan.
constexpr int TC_RR_WARPS = 4;
__global__ void __launch_bounds__(TC_RR_WARPS * 32)
knnTcRerankKernel(ModelView mv, const double* __restrict__ nodes, size_t n, const double* __restrict__ queries,
                  const double* __restrict__ scale, int nqueries, int k, const int* __restrict__ cnt, const int* __restrict__ cand,
                  int gused, int gstride, int capc, int prefixNodes, const unsigned* __restrict__ maxNorm, float* __restrict__ thr,
                  int finalStage, long long* __restrict__ idx, double* __restrict__ dist, unsigned* __restrict__ nflagged,
                  int* __restrict__ flagged, const int32_t* __restrict__ ccOf, int32_t ccFilter, int prefilter) {
  // Error bound of the TF32 accumulator the thresholds are inflated by: eps = 2^epsExp (|a'|^2 + max |b'|^2).  The
  // products are exact in FP32; what is left is the dropped al*bl term and the split residuals (~2^-20 of the norms)
  // and the FP32 accumulation of K = 8 kc slots, whose order and rounding tcgen05 does not specify: at most
  // ~(8 kc + 4) 2^-23 of the largest partial sum (norms + threshold <= 3 (na + nb)): 2^-15 covers kc <= 4 (K <= 32)
  // with a factor > 2, wider models get 2^-14.  Measured on UR5 (kc = 3): 2^-22.
  const int epsExp = (3 * mv.nq + 6 > 32) ? -14 : -15;
  extern __shared__ __align__(16) unsigned char rsm[];
  __shared__ Key lists[TC_RR_WARPS][32];
  __shared__ int sTotal, sKept;
  const Tables T = makeTables(mv.blob, mv);
  double* sw2 = (double*)rsm;
  double* sq = sw2 + mv.njoints;
  int* list = (int*)(sq + mv.nq);  // the query's candidate groups
  int* sJoint = list + TC_LIST;    // per joint: configuration index * 2 + (1 if its distance term is weighted)
  const unsigned FULL = 0xffffffffu;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.x;
  for (int j = threadIdx.x; j < mv.njoints; j += blockDim.x) {
    const double w = T.weights[j];
    sw2[j] = w * w;
    sJoint[j] = T.joints[j].idx_q * 2 + (T.joints[j].type <= HPPB200_JOINT_REVOLUTE_Z ? 1 : 0);
  }
  for (int c = threadIdx.x; c < mv.nq; c += blockDim.x) sq[c] = queries[(size_t)q * mv.nq + c];
  // WeighedDistance(node, query) with the joint records in shared memory: the operation order of
  // weighedDistanceBatch<1, false> (dev_knn.cuh), hence the exact scan's bits
  auto exactDistance = [&](const double* __restrict__ node) {
    double d = 0;
    for (int j = 0; j < mv.njoints; ++j) {
      const int c = sJoint[j] >> 1;
      const double dq = sq[c] - node[c];
      d += (sJoint[j] & 1) ? sw2[j] * (dq * dq) : dq * dq;
    }
    return sqrt(d);
  };
  // 1. the candidate groups, gathered by the lanes of warp 0 in parallel
  if (warp == 0) {
    int total = 0;
    if (cnt == nullptr) {
      total = (prefixNodes + 7) / 8;
      for (int e = lane; e < total; e += 32) list[e] = e;
    } else {
      const int* myCnt = cnt + (size_t)q * gstride;
      int mine = 0;  // lane l owns the lists of CTAs l, l + 32, ...
      bool over = false;
      for (int c = lane; c < gused; c += 32) { const int nc = myCnt[c]; over |= nc > capc; mine += nc; }
      int incl = mine;
      for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += v; }
      total = __shfl_sync(FULL, incl, 31);
      if (__any_sync(FULL, over) || total > TC_LIST) total = -1;
      else {
        int o = incl - mine;
        for (int c = lane; c < gused; c += 32) {
          const int nc = myCnt[c];
          for (int i = 0; i < nc; ++i) list[o++] = cand[(size_t)q * gstride * TC_CAPC + (size_t)c * capc + i];
        }
      }
    }
    if (lane == 0) sTotal = total;
  }
  __syncthreads();
  const int total = sTotal;
  if (total < 0) {
    if (finalStage && threadIdx.x == 0) {
      idx[(size_t)q * k] = -2;
      flagged[atomicAdd(nflagged, 1u)] = q;
    }
    return;  // (not final: the previous threshold stays, it is still an upper bound)
  }
  if (k > 32) {
    // 2'. k > 32: the exact keys not above the current threshold (an upper bound of the k-th squared distance, so no
    // top-k node is dropped) are collected in shared memory and sorted by a bitonic network; first k = the result
    Key* keys = (Key*)(rsm + (((size_t)(mv.njoints + mv.nq) * sizeof(double) + (TC_LIST + mv.njoints) * sizeof(int) + 15) & ~(size_t)15));
    const double bound = cnt == nullptr ? CUDART_INF : (double)thr[q] * (1.0 + 1e-9);
    if (threadIdx.x == 0) sKept = 0;
    __syncthreads();
    for (int e0 = warp * 4; e0 < total; e0 += TC_RR_WARPS * 4) {
      const int e = e0 + (lane >> 3);
      if (e < total) {
        const long long node = (long long)list[e] * 8 + (lane & 7);
        if ((size_t)node < n && (ccFilter < 0 || ccOf[node] == ccFilter)) {
          const double d1 = exactDistance(nodes + (size_t)node * mv.nq);
          if (d1 * d1 <= bound) {
            const int slot = atomicAdd(&sKept, 1);
            if (slot < TC_KEYS) { keys[slot].d = d1; keys[slot].i = node; }
          }
        }
      }
    }
    __syncthreads();
    const int kept = sKept;
    if (kept > TC_KEYS) {
      if (finalStage && threadIdx.x == 0) {
        idx[(size_t)q * k] = -2;
        flagged[atomicAdd(nflagged, 1u)] = q;
      }
      return;
    }
    int P = 64;
    while (P < kept || P < k) P <<= 1;
    for (int x = kept + threadIdx.x; x < P; x += blockDim.x) { keys[x].d = CUDART_INF; keys[x].i = LLONG_MAX; }
    for (int size = 2; size <= P; size <<= 1)
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        __syncthreads();
        for (int t = threadIdx.x; t < P / 2; t += blockDim.x) {
          const int i = (t / stride) * 2 * stride + (t % stride), j = i + stride;
          const Key a = keys[i], b = keys[j];
          if (keyLess(b, a) == ((i & size) == 0)) { keys[i] = b; keys[j] = a; }
        }
      }
    __syncthreads();
    if (finalStage) {
      for (int x = threadIdx.x; x < k; x += blockDim.x) {
        idx[(size_t)q * k + x] = keys[x].i != LLONG_MAX ? keys[x].i : -1;
        dist[(size_t)q * k + x] = keys[x].d;
      }
    } else if (threadIdx.x == 0) {
      float t = TC_THR_MAX;
      if (keys[k - 1].i != LLONG_MAX) {
        double na = 0;
        for (int c = 0; c < mv.nq; ++c) { const double x = scale[c] * sq[c]; na = fma(x, x, na); }
        const double eps = ldexp(na + (double)__uint_as_float(*maxNorm), epsExp) + 1e-30;
        t = fminf(TC_THR_MAX, __double2float_ru(keys[k - 1].d * keys[k - 1].d * (1.0 + 1e-6) + eps));
      }
      if (cnt == nullptr || t < thr[q]) thr[q] = t;
    }
    return;
  }
  // 2. exact distances: 4 groups of 8 nodes per warp and trip
  Key acc = {CUDART_INF, LLONG_MAX}, ak = acc;
  auto offer = [&](const Key c) {
    unsigned m = __ballot_sync(FULL, keyLess(c, ak));
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1u;
      const Key x = {__shfl_sync(FULL, c.d, src), __shfl_sync(FULL, c.i, src)};
      if (keyLess(x, ak)) topkInsert(acc, ak, x, k, lane);
    }
  };
  const double bound2 = (cnt == nullptr || !prefilter) ? CUDART_INF : (double)thr[q] * (1.0 + 1e-9);
  for (int e0 = warp * 4; e0 < total; e0 += TC_RR_WARPS * 4) {
    const int e = e0 + (lane >> 3);
    Key c = {CUDART_INF, LLONG_MAX};
    if (e < total) {
      const long long node = (long long)list[e] * 8 + (lane & 7);
      if ((size_t)node < n && (ccFilter < 0 || ccOf[node] == ccFilter)) {
        const double d1 = exactDistance(nodes + (size_t)node * mv.nq);
        if (d1 * d1 <= bound2) { c.d = d1; c.i = node; }  // above the threshold (an upper bound of the k-th): never in the top k
      }
    }
    offer(c);
  }
  lists[warp][lane] = acc;
  __syncthreads();
  if (warp != 0) return;
  for (int w = 1; w < TC_RR_WARPS; ++w) {
    Key c = lists[w][lane];
    if (lane >= k) { c.d = CUDART_INF; c.i = LLONG_MAX; }
    offer(c);
  }
  if (finalStage) {
    if (lane < k) {
      idx[(size_t)q * k + lane] = acc.i != LLONG_MAX ? acc.i : -1;
      dist[(size_t)q * k + lane] = acc.d;
    }
  } else if (lane == 0) {
    float t = TC_THR_MAX;
    if (ak.i != LLONG_MAX) {
      double na = 0;
      for (int c = 0; c < mv.nq; ++c) { const double x = scale[c] * sq[c]; na = fma(x, x, na); }
      const double eps = ldexp(na + (double)__uint_as_float(*maxNorm), epsExp) + 1e-30;
      t = fminf(TC_THR_MAX, __double2float_ru(ak.d * ak.d * (1.0 + 1e-6) + eps));
    }
    if (cnt == nullptr || t < thr[q]) thr[q] = t;
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
static size_t tcSmemBytes(int nqt, int kc, int bstages) {
  return 512 + (size_t)nqt * TC_M * kc * 32 + (size_t)bstages * TC_N * kc * 32 + (size_t)nqt * TC_M * sizeof(int);
}
static int tcMaxQueryTiles(int kc) {
  int nqt = TC_MAX_QT;
  while (nqt > 0 && tcSmemBytes(nqt, kc, 2) > 220 * 1024) --nqt;
  return nqt;
}
static int tcStages(int nqt, int kc) {
  int st = TC_BSTAGES;
  while (st > 2 && tcSmemBytes(nqt, kc, st) > 220 * 1024) --st;
  return st;
}

void knnTcInit(hppb200_knn* nn) {
  if (const char* v = std::getenv("HPPB200_KNN_TC")) nn->tc_enabled = std::atoi(v) != 0;
  if (const char* v = std::getenv("HPPB200_KNN_TC_MIN")) nn->tc_min_queries = std::max(1, std::atoi(v));
  if (const char* v = std::getenv("HPPB200_KNN_TC_PREFILTER")) nn->tc_prefilter = std::atoi(v) != 0;
  nn->tc_ratio = TC_RATIO;
  if (const char* v = std::getenv("HPPB200_KNN_TC_RATIO")) nn->tc_ratio = std::max(2.0, std::atof(v));
  nn->tc_kc = 0;
  if (nn->ff) return;
  const int nq = nn->model->nq;
  const int kc = (3 * nq + 6 + 7) / 8;
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
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(knnTcFilterKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(knnTcFilterKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  HPPB200_CUDA_CHECK(cudaEventCreate(&nn->tc_ev0));
  HPPB200_CUDA_CHECK(cudaEventCreate(&nn->tc_ev1));
}

void knnTcDestroy(hppb200_knn* nn) {
  if (nn->d_tc_nodes) cudaFree(nn->d_tc_nodes);
  nn->d_tc_nodes = nullptr;
  if (nn->tc_ev0) cudaEventDestroy(nn->tc_ev0);
  if (nn->tc_ev1) cudaEventDestroy(nn->tc_ev1);
  nn->tc_ev0 = nn->tc_ev1 = nullptr;
}

bool knnTcEligible(const hppb200_knn* nn, size_t nqueries, int k, int32_t cc) {
  (void)cc;  // a component filter is applied by the exact re-rank; sparse components end in the exact scan
  return nn->tc_enabled && nn->tc_kc > 0 && k >= 1 && k <= TC_MAX_K && nn->n >= TC_MIN_NODES &&
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
                                                                                    rowsEnd, nn->n, nn->d_tc_nodes, nn->d_tc_misc.p, nullptr);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  nn->tc_n = nn->n;
}

// one filter pass of `nqtTotal` query tiles over the first `prefixNodes` nodes; returns the grid used
static int tcLaunchFilter(hppb200_knn* nn, size_t prefixNodes, int nqtTotal, float* dump, int swapOffsets) {
  const int kc = nn->tc_kc;
  const int perLaunch = tcMaxQueryTiles(kc);
  const unsigned ntiles = (unsigned)((prefixNodes + TC_N - 1) / TC_N);
  const unsigned grid = std::min<unsigned>((unsigned)nn->num_sms, ntiles);
  for (int qt0 = 0; qt0 < nqtTotal; qt0 += perLaunch) {
    TcFilterArgs a;
    a.nqt = std::min(perLaunch, nqtTotal - qt0);
    a.kc = kc;
    a.bstages = tcStages(a.nqt, kc);
    a.gstride = nn->num_sms;
    a.capc = TC_CAPC * (nn->num_sms / (int)grid);  // few CTAs (small prefix): longer lists each
    a.btiles = nn->d_tc_nodes;
    a.atiles = nn->d_tc_queries.p + (size_t)qt0 * TC_M * 8 * kc;
    a.cnt = nn->d_tc_cnt.p + (size_t)qt0 * TC_M * a.gstride;
    a.cand = nn->d_tc_cand.p + (size_t)qt0 * TC_M * a.gstride * TC_CAPC;
    a.dump = qt0 == 0 ? dump : nullptr;
    a.ntiles = ntiles;
    a.swapOffsets = swapOffsets;
    { const char* v = std::getenv("HPPB200_KNN_TC_ABLATE"); a.ablate = v ? std::atoi(v) : 0; }
    a.prof = nullptr;
    static const bool wantProf = std::getenv("HPPB200_KNN_TC_PROF") != nullptr;  // timing experiments: per-CTA wait cycles to stderr
    DeviceBuffer<long long> prof;
    if (wantProf) { prof.reserve((size_t)grid * 8); a.prof = prof.p; }
    if (wantProf) knnTcFilterKernel<true><<<grid, TC_THREADS, tcSmemBytes(a.nqt, kc, a.bstages), nn->stream>>>(a);
    else knnTcFilterKernel<false><<<grid, TC_THREADS, tcSmemBytes(a.nqt, kc, a.bstages), nn->stream>>>(a);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    if (wantProf) {
      std::vector<long long> h((size_t)grid * 8);
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
      HPPB200_CUDA_CHECK(cudaMemcpy(h.data(), prof.p, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
      double m[8] = {0};
      for (unsigned c = 0; c < grid; ++c) for (int i = 0; i < 8; ++i) m[i] += (double)h[(size_t)c * 8 + i] / grid;
      std::fprintf(stderr, "knnTcFilter prof: tiles %u nqt %d kc %d stages %d | mma thread: wait B %.0f wait acc %.0f issue %.0f total %.0f | "
                   "epilogue warp 0: wait acc %.0f drain %.0f total %.0f uses %.0f (cycles, mean over %u CTAs)\n",
                   ntiles, a.nqt, kc, a.bstages, m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], grid);
    }
  }
  return (int)grid;
}

static void tcReserve(hppb200_knn* nn, int nqtTotal) {
  const size_t rows = (size_t)nqtTotal * TC_M;
  nn->d_tc_queries.reserve(rows * 8 * nn->tc_kc);
  nn->d_tc_thr.reserve(rows);
  nn->d_tc_cnt.reserve(rows * nn->num_sms);
  nn->d_tc_cand.reserve(rows * nn->num_sms * TC_CAPC);
  nn->d_tc_flagged.reserve(rows);
}
// query operand tiles with the current thresholds folded in (thr == null: 0)
static void tcPrepQueries(hppb200_knn* nn, const double* d_q, size_t nqueries, int nqtTotal, const float* thr) {
  const size_t rows = (size_t)nqtTotal * TC_M;
  knnTcPrepKernel<<<(unsigned)((rows + 127) / 128), 128, 0, nn->stream>>>(d_q, nn->d_tc_scale.p, nn->model->nq, nn->tc_kc, 0, TC_M, 0, rows,
                                                                          nqueries, nn->d_tc_queries.p, nn->d_tc_misc.p, thr);
  HPPB200_CUDA_CHECK(cudaGetLastError());
}

static void tcRunChunk(hppb200_knn* nn, const double* d_q, size_t nqueries, int k, int32_t cc, long long* d_i, double* d_d) {
  const int nq = nn->model->nq;
  const int nqtTotal = (int)((nqueries + TC_M - 1) / TC_M);
  tcReserve(nn, nqtTotal);
  const size_t rerankSmem = ((size_t)nn->model->njoints + (size_t)nq) * sizeof(double) + (TC_LIST + (size_t)nn->model->njoints) * sizeof(int) + 16 +
                            (k > 32 ? TC_KEYS * sizeof(Key) : 0);
  const unsigned rerankGrid = (unsigned)nqueries;
  HPPB200_CUDA_CHECK(cudaMemsetAsync(nn->d_tc_misc.p + 1, 0, sizeof(unsigned), nn->stream));
  // bootstrap: exact k-th distance over the first TC_PREFIX0 nodes -> first thresholds
  knnTcRerankKernel<<<rerankGrid, TC_RR_WARPS * 32, rerankSmem, nn->stream>>>(nn->view, nn->d_nodes, nn->n, d_q, nn->d_tc_scale.p, (int)nqueries, k, nullptr,
                                                                 nullptr, 0, 0, 0, (int)TC_PREFIX0, nn->d_tc_misc.p, nn->d_tc_thr.p, 0, d_i, d_d,
                                                                 nn->d_tc_misc.p + 1, nn->d_tc_flagged.p, nn->d_cc, cc, nn->tc_prefilter ? 1 : 0);
  // prefixes growing geometrically (<= TC_RATIO per stage) up to the whole roadmap
  const double ratio = (double)nn->n / (double)TC_PREFIX0;
  // (large k: about 512 candidate groups per query and stage at most, so that the per-CTA lists do not overflow)
  const double stageRatio = std::max(2.0, std::min(nn->tc_ratio, 512.0 / (double)k));
  const int nstages = std::max(1, (int)std::ceil(std::log(ratio) / std::log(stageRatio) - 1e-9));
  for (int s = 1; s <= nstages; ++s) {
    size_t prefix = nn->n;
    if (s < nstages) {
      prefix = (size_t)((double)TC_PREFIX0 * std::pow(ratio, (double)s / nstages));
      prefix = std::min(nn->n, (prefix + TC_N - 1) / TC_N * TC_N);
    }
    tcPrepQueries(nn, d_q, nqueries, nqtTotal, nn->d_tc_thr.p);
    if (s == nstages) HPPB200_CUDA_CHECK(cudaEventRecord(nn->tc_ev0, nn->stream));  // the whole-roadmap pass, timed
    const int gused = tcLaunchFilter(nn, prefix, nqtTotal, nullptr, 0);
    if (s == nstages) HPPB200_CUDA_CHECK(cudaEventRecord(nn->tc_ev1, nn->stream));
    knnTcRerankKernel<<<rerankGrid, TC_RR_WARPS * 32, rerankSmem, nn->stream>>>(nn->view, nn->d_nodes, nn->n, d_q, nn->d_tc_scale.p, (int)nqueries, k,
                                                                   nn->d_tc_cnt.p, nn->d_tc_cand.p, gused, nn->num_sms, TC_CAPC * (nn->num_sms / gused), 0,
                                                                   nn->d_tc_misc.p,
                                                                   nn->d_tc_thr.p, s == nstages ? 1 : 0, d_i, d_d, nn->d_tc_misc.p + 1,
                                                                   nn->d_tc_flagged.p, nn->d_cc, cc, nn->tc_prefilter ? 1 : 0);
    HPPB200_CUDA_CHECK(cudaGetLastError());
  }
  // queries with an overflowed list in the last stage: exact scan
  unsigned nflagged = 0;
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&nflagged, nn->d_tc_misc.p + 1, sizeof(unsigned), cudaMemcpyDeviceToHost, nn->stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
  float fms = 0;
  HPPB200_CUDA_CHECK(cudaEventElapsedTime(&fms, nn->tc_ev0, nn->tc_ev1));
  nn->tc_last_filter_ms += fms;
  nn->tc_last_filter_pairs += (double)nqtTotal * TC_M * (double)((nn->n + TC_N - 1) / TC_N * TC_N);
  nn->tc_last_stages = nstages;
  nn->tc_last_overflow += nflagged;
  nn->tc_last_rows = (size_t)nqtTotal * TC_M;
  nn->tc_last_gused = (int)std::min<size_t>((size_t)nn->num_sms, (nn->n + TC_N - 1) / TC_N);
  if (nflagged) {
    DeviceBuffer<double> gq, gd;
    DeviceBuffer<long long> gi;
    gq.reserve((size_t)nflagged * nq);
    gd.reserve((size_t)nflagged * k);
    gi.reserve((size_t)nflagged * k);
    knnTcGatherQueriesKernel<<<(nflagged * (unsigned)nq + 255) / 256, 256, 0, nn->stream>>>(d_q, nn->d_tc_flagged.p, (int)nflagged, nq, gq.p);
    knnExactRun(nn, gq.p, nflagged, k, cc, nn->n, gi.p, gd.p);
    knnTcScatterResultsKernel<<<(nflagged * (unsigned)k + 255) / 256, 256, 0, nn->stream>>>(gi.p, gd.p, nn->d_tc_flagged.p, (int)nflagged, k,
                                                                                           d_i, d_d);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));
  }
}

void knnTcRun(hppb200_knn* nn, const double* d_q, size_t nqueries, int k, int32_t cc, long long* d_i, double* d_d) {
  tcSyncNodeOperand(nn);
  nn->tc_last_used = true;
  nn->tc_last_overflow = 0;
  nn->tc_last_filter_ms = 0;
  nn->tc_last_filter_pairs = 0;
  const size_t chunk = 8192;  // bounds the candidate buffers (9.5 KB per query)
  for (size_t q0 = 0; q0 < nqueries; q0 += chunk) {
    const size_t nqc = std::min(chunk, nqueries - q0);
    tcRunChunk(nn, d_q + q0 * (size_t)nn->model->nq, nqc, k, cc, d_i + q0 * (size_t)k, d_d + q0 * (size_t)k);
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
      std::vector<int> h(nn->tc_last_rows * (size_t)nn->num_sms);  // (of the last chunk of 8192 queries)
      HPPB200_CUDA_CHECK(cudaMemcpy(h.data(), nn->d_tc_cnt.p, h.size() * sizeof(int), cudaMemcpyDeviceToHost));
      for (size_t r = 0; r < nn->tc_last_rows; ++r)
        for (int c = 0; c < nn->tc_last_gused; ++c) out[3] += h[r * (size_t)nn->num_sms + c];
    }
    return HPPB200_OK;
  } catch (const std::exception& e) {
    setLastError(e.what());
    return std::strncmp(e.what(), "CUDA", 4) == 0 ? HPPB200_ERR_CUDA : HPPB200_ERR_INVALID_ARGUMENT;
  }
}

// ------------------------------------------------------------------------------------------------
// Measured TMEM -> register bandwidth: the roofline denominator of knnTcFilterKernel, whose output (one FP32 word per
// (query, node) pair) has to leave tensor memory through tcgen05.ld.  One CTA per SM allocates the whole tensor memory
// (512 columns); 16 warps (warp w reads TMEM lanes 32 (w % 4) .., columns 128 (w / 4) ..: every warp can only reach its own
// lane quarter) stream it with the filter's own access shape (tcgen05.ld.32x32b.x32, two loads in flight per warp) and OR
// the words so that nothing is optimised away.  Bytes = loads x 32 lanes x 32 columns x 4 B.
__global__ void __launch_bounds__(512, 1) tmemReadPeakKernel(int iters, unsigned* sink) {
  __shared__ uint32_t tmemPtr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(&tmemPtr)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcFenceBefore();
  __syncthreads();
  tcFenceAfter();
  const uint32_t base = tmemPtr + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(128 * (warp >> 2));
  uint32_t a[32], b[32];
  unsigned acc = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int c = 0; c < 128; c += 64) {
      tcLoad32(base + c, a);
      tcLoad32(base + c + 32, b);
      tcWaitLoad(a);
      tcWaitLoad(b);
#pragma unroll
      for (int i = 0; i < 32; ++i) acc |= a[i] | b[i];
    }
  }
  if (acc == 0x12345u) sink[0] = acc;  // (never true in practice: keeps the loads alive)
  tcFenceBefore();
  __syncthreads();
  if (warp == 0) {
    tcFenceAfter();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemPtr), "r"(512u) : "memory");
  }
}

int hppb200_measure_tmem_read_peak(int device, double* tb_per_s) {
  try {
    if (!tb_per_s) throw std::invalid_argument("null argument");
    if (hppb200_device_count() == 0) throw std::runtime_error("no CUDA device visible: the engine has no CPU fallback");
    HPPB200_CUDA_CHECK(cudaSetDevice(device));
    cudaDeviceProp prop;
    HPPB200_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    unsigned* sink = nullptr;
    HPPB200_CUDA_CHECK(cudaMalloc((void**)&sink, sizeof(unsigned)));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 4096;
    double best = 0;
    for (int rep = 0; rep < 5; ++rep) {
      cudaEventRecord(e0);
      tmemReadPeakKernel<<<prop.multiProcessorCount, 512>>>(iters, sink);
      cudaEventRecord(e1);
      HPPB200_CUDA_CHECK(cudaEventSynchronize(e1));
      HPPB200_CUDA_CHECK(cudaGetLastError());
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double bytes = (double)prop.multiProcessorCount * 16 /* warps */ * iters * 4 /* loads per turn */ * 32 * 32 * 4;
      if (rep > 0) best = std::max(best, bytes / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *tb_per_s = best;
    return HPPB200_OK;
  } catch (std::invalid_argument& e) { setLastError(e.what()); return HPPB200_ERR_INVALID_ARGUMENT; }
  catch (std::exception& e) { setLastError(e.what()); return HPPB200_ERR_CUDA; }
}

int hppb200_knn_tc_filter_time(hppb200_knn* nn, double* ms, double* pairs) {
  if (!nn || !ms || !pairs) { setLastError("null argument"); return HPPB200_ERR_INVALID_ARGUMENT; }
  *ms = nn->tc_last_used ? nn->tc_last_filter_ms : 0.0;
  *pairs = nn->tc_last_used ? nn->tc_last_filter_pairs : 0.0;
  return HPPB200_OK;
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
    tcReserve(nn, 1);
    tcPrepQueries(nn, nn->d_queries.p, nqueries, 1, nullptr);  // threshold 0: the accumulator is the squared distance
    DeviceBuffer<float> dump;
    dump.reserve((size_t)TC_M * 2 * TC_N);
    HPPB200_CUDA_CHECK(cudaMemsetAsync(dump.p, 0, (size_t)TC_M * 2 * TC_N * sizeof(float), nn->stream));
    tcLaunchFilter(nn, std::min<size_t>(nn->n, 2 * TC_N), 1, dump.p, swap_offsets);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(out, dump.p, (size_t)TC_M * 2 * TC_N * sizeof(float), cudaMemcpyDeviceToHost, nn->stream));
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
