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
#include <cstdio>

#include "dev_model.cuh"

namespace hppb200 {

constexpr int CFG_THREADS = 256;
constexpr int MODE_VALIDATE = 0, MODE_DISTANCES = 1, MODE_PLACEMENTS = 2;

struct ConfigSmemLayout {
  uint32_t off_blob, off_store, off_first, off_bound, total;
};

static ConfigSmemLayout configSmemLayout(const ModelView& mv, bool tree, uint32_t CFG_PER_BLOCK) {
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

// CFG_PER_BLOCK configurations per CTA: 64, or 32 when the placements of a larger model would not fit
template <bool TREE, int CFG_PER_BLOCK>
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
      if (boundHit[tid] < 0) forwardKinematics<TREE>(T, mv, ConfigQ{qc}, store + tid, CFG_PER_BLOCK);
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

// =====================================================================================
// Production path for MODE_VALIDATE: two kernels.
//
//  K1 configPassKernel — one THREAD per configuration, uniform work:
//       joint bounds, FK in registers (sparse joint transform), geometry placements written
//       once to a global scratch (SoA over configurations, coalesced; read back only by the
//       rare GJK jobs) and, per geometry, the 6 doubles the cheap tests need (core centre +
//       capsule axis) kept in shared memory.  Then the pair loop IN REFERENCE ORDER with the
//       reference's early exit: closed-form kinds (sphere/capsule) are decided exactly;
//       GJK kinds first get a conservative cull test (squared core distance vs. inflation
//       radii + margin + 1e-9 slack, then a box-face separating-axis bound): a culled pair
//       provably does not collide, so flags stay identical to the reference; a pair that
//       survives is queued as a job (configuration, pair).  ~99 % of the reference's GJK
//       calls are culled on the UR5 + boxes workload.
//  K2 gjkJobsKernel — one thread per surviving job, full warps of GJK; a colliding job does
//       atomicMin on the configuration's first-hit index (unsigned: -1 = none = max).
// The minimum colliding index over exact tests equals the reference's first hit.
// =====================================================================================
constexpr int K1_THREADS = 128;
constexpr int K1_JOBS_PER_BLOCK = 4 * K1_THREADS;  // smem staging; overflow goes straight to global

struct ValidateArgs {
  const double* q;
  size_t n;
  uint32_t flags;
  uint8_t* valid;
  uint32_t* first;  // int32 viewed as unsigned (0xffffffff = -1)
  uint8_t* reason;
  double* oMg;             // [ngeoms*12][n] scratch
  unsigned long long* jobs;  // (config << 32 | pair)
  unsigned long long* njobs;
  unsigned long long jobs_cap;
  // configPassKernel stages the blob WITHOUT the continuous-validation tables (body pairs, their geometry pairs,
  // velocity coefficients): [0, blob_head) and [blob_tail_off, blob_tail_off + blob_tail_bytes) back to back; the
  // kernel's ModelView has the offsets of the tables behind the gap moved down accordingly (configView below)
  uint32_t blob_head, blob_tail_off, blob_tail_bytes;
  int force_exact;         // SHADOW: every pair from the FP64 placements (HPPB200_CFG_SHADOW=2, tests)
  double hmax, reach_max;  // SHADOW slack: largest capsule half-length of a robot geometry, largest cull reach of a pair
  // DMIN instances (first-sample screen of the edge path): [n] how close this configuration is to a collision: the smallest of the
  // distances of the pairs that were evaluated (closed forms here, GJK jobs in gjkJobsKernel) and of the separation bounds of the
  // pairs that were culled (a culled pair is provably apart, possibly by a fraction of a millimetre), as the bits of a non-negative
  // double.  A scheduling hint only (edges that start close to an obstacle take the most samples and are started first); it
  // reaches no result.
  unsigned long long* dmin;
};

// Two TMA bulk copies on one mbarrier (see stageBlobTMA): the blob without the gap.
__device__ __forceinline__ void stageBlobSplitTMA(unsigned char* sdst, const unsigned char* gsrc, uint32_t head, uint32_t tailOff,
                                                  uint32_t tailBytes, uint64_t* bar) {
  const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(bar);
  const uint32_t dst_s = (uint32_t)__cvta_generic_to_shared(sdst);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(head + tailBytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s), "l"(gsrc),
                 "r"(head), "r"(bar_s)
                 : "memory");
    if (tailBytes)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s + head),
                   "l"(gsrc + tailOff), "r"(tailBytes), "r"(bar_s)
                   : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar_s), "r"(0u)
        : "memory");
  }
}

// Sparse pinocchio joint transform: liMi = P * M_joint(q), exploiting the zeros/ones of M.
// Bit-identical to the dense product up to the sign of zeros.
__device__ __forceinline__ T3 jointLocalSparse(const DevJoint& J, double qv) {
  T3 P = load12(J.P);
  T3 R = P;
  if (J.type <= HPPB200_JOINT_REVOLUTE_Z) {
    double s, c;
    sincos_portable(qv, &s, &c);
    const double ms = -s;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double p0 = P.r[3 * i], p1 = P.r[3 * i + 1], p2 = P.r[3 * i + 2];
      if (J.type == HPPB200_JOINT_REVOLUTE_Z) { R.r[3 * i] = fma(p1, s, p0 * c); R.r[3 * i + 1] = fma(p1, c, p0 * ms); }
      else if (J.type == HPPB200_JOINT_REVOLUTE_Y) { R.r[3 * i] = fma(p2, ms, p0 * c); R.r[3 * i + 2] = fma(p2, c, p0 * s); }
      else { R.r[3 * i + 1] = fma(p2, s, p1 * c); R.r[3 * i + 2] = fma(p2, c, p1 * ms); }
    }
  } else {
    const int a = J.type - HPPB200_JOINT_PRISMATIC_X;
    // R unchanged; t = P.R[:,a] * q + P.t  (dense form: dot(row, (..q..)) + t)
    R.t = mk(P.r[a] * qv + P.t.x, P.r[3 + a] * qv + P.t.y, P.r[6 + a] * qv + P.t.z);
  }
  return R;
}

__device__ __forceinline__ void pushJob(unsigned long long job, unsigned long long* sjobs, int* scount, const ValidateArgs& A) {
  const int slot = atomicAdd(scount, 1);
  if (slot < K1_JOBS_PER_BLOCK) { sjobs[slot] = job; return; }
  const unsigned long long g = atomicAdd(A.njobs, 1ull);  // staging full (rare): straight to global
  if (g < A.jobs_cap) A.jobs[g] = job;
}

// SHADOW (the production form): the bounding cores of the robot geometries are kept in shared memory as the HIGH 32-bit
// words of their FP64 coordinates (sign, exponent, 20 mantissa bits; rebuilt with a zero low word: no conversion
// instruction), 24 instead of 48 bytes per geometry and thread, so that 5 CTAs instead of 3 fit an SM (the kernel is
// latency-bound: 0.92 -> see DESIGN.md 3.1).  Decisions stay EXACT:
//  * truncation moves a coordinate by < 2^-20 |x|, a core point (centre +- h axis) by < sqrt(3) 2^-20 (maxc + h), and
//    core distances are 1-Lipschitz in their points, so |d~ - d| < 2 sqrt(3) 2^-20 (maxc + hmax) < 2^-18 (maxc + hmax);
//    the point-segment form assumes a unit axis, |u~| is short by < 2^-19, which moves its foot point by < 2^-19 h.
//    s = 2^-17 (maxc + hmax) + 1e-9 covers both twice over (maxc = largest |coordinate| of this configuration's cores);
//  * closed forms: d~ - margin > threshold + s  => no collision;  <= threshold - s  => collision (the exact test says
//    the same in both cases);  in between (|d - margin| < 2 s, ~1 pair in 10^4) the pair is evaluated again from the
//    FP64 placements in the global scratch, with the operations of the FP64 path;
//  * conservative culls only get the slack added (fewer culls, never a wrong one): core distance^2 > cull2 +
//    s (2 reach_max + s), box-face bound > 1e-9 + s;
//  * a configuration with NaN / infinite cores or maxc > 64 takes the exact path for every pair.
#ifndef K1_MINB
#define K1_MINB 6
#endif
__device__ __forceinline__ double hiWord(uint32_t h) { return __hiloint2double((int)h, 0); }

// HOSTQ: A.q is PINNED HOST memory (zero-copy form of the host-pointer path): thread 0 pulls the CTA's 128 configurations
// (one contiguous block) over PCIe with one TMA bulk copy into shared memory while the blob is staged; the copy engine, its
// slice-sized latency and the device staging buffer drop out, and the transfer overlaps the FP64 work CTA by CTA.
template <bool TREE, bool SHADOW, bool HOSTQ, bool DMIN = false>
__global__ void __launch_bounds__(K1_THREADS, SHADOW ? K1_MINB : 1)
configPassKernel(ModelView mv, ValidateArgs A, DeviceStats* stats) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  int* scount = (int*)(smem + 8);
  unsigned char* sblob = smem + 16;
  double* cstore = (double*)(sblob + mv.blob_bytes);                         // [ngeoms*6][K1_THREADS] f64, or u32 (SHADOW)
  uint32_t* cstore32 = (uint32_t*)cstore;
  unsigned long long* sjobs = (unsigned long long*)((unsigned char*)cstore + (size_t)mv.ngeoms * 6 * K1_THREADS * (SHADOW ? 4 : 8));
  __shared__ unsigned long long sbase;
  __shared__ __align__(8) uint64_t qbar;
  const int tid = threadIdx.x;
  double* sq = (double*)sjobs;  // HOSTQ: [K1_THREADS][nq], in the job staging area (used only after phase 1)
  bool qStaged = false;
  if (HOSTQ) {
    const size_t base = (size_t)blockIdx.x * K1_THREADS;
    const size_t cnt = A.n - base < (size_t)K1_THREADS ? A.n - base : (size_t)K1_THREADS;
    const uint32_t bytes = (uint32_t)(cnt * (size_t)mv.nq * sizeof(double));
    // (a partial CTA with an odd number of doubles, or a misaligned block, reads its rows directly)
    qStaged = (bytes & 15u) == 0u && ((uintptr_t)(A.q + base * (size_t)mv.nq) & 15u) == 0u;
    if (qStaged && tid == 0) {
      const uint32_t qb = (uint32_t)__cvta_generic_to_shared(&qbar);
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(qb) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(qb), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       (uint32_t)__cvta_generic_to_shared(sq)),
                   "l"(A.q + base * (size_t)mv.nq), "r"(bytes), "r"(qb)
                   : "memory");
    }
  }
  stageBlobSplitTMA(sblob, mv.blob, A.blob_head, A.blob_tail_off, A.blob_tail_bytes, bar);
  const Tables T = makeTables(sblob, mv);
  if (tid == 0) *scount = 0;
  __syncthreads();
  if (HOSTQ && qStaged) {
    const uint32_t qb = (uint32_t)__cvta_generic_to_shared(&qbar);
    uint32_t done = 0;
    while (!done) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(qb), "r"(0u)
          : "memory");
    }
  }
  const size_t i = (size_t)blockIdx.x * K1_THREADS + tid;
  unsigned tests = 0;
  // phase 1: joint bounds + FK (the only reader of the configuration); phase 2: the pair loop.  HOSTQ separates them
  // with a barrier because the staged rows share their shared memory with the job staging of phase 2.
  bool live = false;
  double* cs = cstore + tid;
  uint32_t* cs32 = cstore32 + tid;
  double maxc = 0.0;
  bool exactAll = !SHADOW || A.force_exact != 0;
  if (i < A.n) {
    const double* qc = (HOSTQ && qStaged) ? sq + (size_t)tid * mv.nq : A.q + i * (size_t)mv.nq;
    int bad = -1;
    if (A.flags & HPPB200_CHECK_JOINT_BOUNDS) {
      for (int k = 0; k < mv.nq && bad < 0; ++k) {
        const double lo = T.lower[k], hi = T.upper[k];
        if (lo > hi) continue;
        if (qc[k] < lo || hi < qc[k]) bad = k;
      }
    }
    if (bad >= 0) {
      A.valid[i] = 0; A.first[i] = (uint32_t)bad; A.reason[i] = HPPB200_REASON_JOINT_BOUND;
    } else {
      // ---- FK: placements to global scratch, compact cores to shared memory ----
      live = true;
      auto emit = [&](int g, const T3& M) {
        store12s(A.oMg + (size_t)g * 12 * A.n + i, A.n, M);
        const DevGeom& G = T.geoms[g];
        V3 c = M.t;
        if (G.shape.type >= SH_BOX) c = rmul(M, mk(G.bound.bc[0], G.bound.bc[1], G.bound.bc[2])) + M.t;
        if (SHADOW) {
          cs32[(g * 6 + 0) * K1_THREADS] = (uint32_t)__double2hiint(c.x);
          cs32[(g * 6 + 1) * K1_THREADS] = (uint32_t)__double2hiint(c.y);
          cs32[(g * 6 + 2) * K1_THREADS] = (uint32_t)__double2hiint(c.z);
          cs32[(g * 6 + 3) * K1_THREADS] = (uint32_t)__double2hiint(M.r[2]);
          cs32[(g * 6 + 4) * K1_THREADS] = (uint32_t)__double2hiint(M.r[5]);
          cs32[(g * 6 + 5) * K1_THREADS] = (uint32_t)__double2hiint(M.r[8]);
          maxc = fmax(maxc, fmax(fabs(c.x), fmax(fabs(c.y), fabs(c.z))));
          const double sum = ((c.x + c.y) + c.z) + ((M.r[2] + M.r[5]) + M.r[8]);
          if (!(fabs(sum) < CUDART_INF)) exactAll = true;  // NaN or infinite somewhere
        } else {
          cs[(g * 6 + 0) * K1_THREADS] = c.x; cs[(g * 6 + 1) * K1_THREADS] = c.y; cs[(g * 6 + 2) * K1_THREADS] = c.z;
          cs[(g * 6 + 3) * K1_THREADS] = M.r[2]; cs[(g * 6 + 4) * K1_THREADS] = M.r[5]; cs[(g * 6 + 5) * K1_THREADS] = M.r[8];
        }
      };
      for (int g = 0; g < mv.ngeoms; ++g)
        if (T.geoms[g].joint == 0) emit(g, load12(T.geoms[g].P));
      T3 cur = identityT();
      for (int j = 1; j <= mv.njoints; ++j) {
        const DevJoint& J = T.joints[j - 1];
        T3 li = J.type == HPPB200_JOINT_FREEFLYER ? jointLocal(J, ConfigQ{qc}) : jointLocalSparse(J, qc[J.idx_q]);
        if (J.parent == 0) cur = li;
        else if (!TREE || J.parent == j - 1) cur = compose(cur, li);
        else cur = compose(load12s(A.oMg + (size_t)(mv.ngeoms + J.parent - 1) * 12 * A.n + i, A.n), li);
        if (TREE) store12s(A.oMg + (size_t)(mv.ngeoms + j - 1) * 12 * A.n + i, A.n, cur);
        for (int g = 0; g < mv.ngeoms; ++g)
          if (T.geoms[g].joint == j) emit(g, compose(cur, load12(T.geoms[g].P)));
      }
    }
  }
  if (HOSTQ) __syncthreads();
  {
    if (live) {
      if (SHADOW && !(maxc <= 64.0)) exactAll = true;
      const double slack = SHADOW ? fma(0x1p-17, maxc + A.hmax, 1e-9) : 0.0;
      const double cullSlack = SHADOW ? slack * fma(2.0, A.reach_max, slack) * (1.0 + 1e-9) : 0.0;
      // ---- pair loop in reference order, early exit at the first definite hit ----
      auto core = [&](int ref, bool ex, V3& c, V3& u) {
        if (ref >= 0) {
          if (!SHADOW) {
            c = mk(cs[(ref * 6 + 0) * K1_THREADS], cs[(ref * 6 + 1) * K1_THREADS], cs[(ref * 6 + 2) * K1_THREADS]);
            u = mk(cs[(ref * 6 + 3) * K1_THREADS], cs[(ref * 6 + 4) * K1_THREADS], cs[(ref * 6 + 5) * K1_THREADS]);
          } else if (!ex) {
            c = mk(hiWord(cs32[(ref * 6 + 0) * K1_THREADS]), hiWord(cs32[(ref * 6 + 1) * K1_THREADS]), hiWord(cs32[(ref * 6 + 2) * K1_THREADS]));
            u = mk(hiWord(cs32[(ref * 6 + 3) * K1_THREADS]), hiWord(cs32[(ref * 6 + 4) * K1_THREADS]), hiWord(cs32[(ref * 6 + 5) * K1_THREADS]));
          } else {  // the FP64 placement this thread wrote to the global scratch: the values the FP64 store would hold
            const double* pm = A.oMg + (size_t)ref * 12 * A.n + i;
            const DevGeom& G = T.geoms[ref];
            if (G.shape.type >= SH_BOX) {
              const T3 M = load12s(pm, A.n);
              c = rmul(M, mk(G.bound.bc[0], G.bound.bc[1], G.bound.bc[2])) + M.t;
              u = mk(M.r[2], M.r[5], M.r[8]);
            } else {
              c = mk(pm[9 * A.n], pm[10 * A.n], pm[11 * A.n]);
              u = mk(pm[2 * A.n], pm[5 * A.n], pm[8 * A.n]);
            }
          }
        } else {
          const DevObstacle& O = T.obstacles[-(ref + 1)];
          c = mk(O.wc[0], O.wc[1], O.wc[2]);
          u = mk(O.P[2], O.P[5], O.P[8]);
        }
      };
      double dminLocal = CUDART_INF;  // DMIN: smallest distance (evaluated pairs) or separation bound (culled pairs) of this configuration
      // box-face separating axes: a lower bound of the distance of geometry a's core to the obstacle box b (true: culled)
      auto boxFaceCull = [&](int p, bool ex) -> bool {
        const DevPair pr = T.cpairs[p];
        const DevShape& sa = *(const DevShape*)(pr.a >= 0 ? &T.geoms[pr.a].shape : &T.obstacles[-(pr.a + 1)].shape);
        const DevObstacle& O = T.obstacles[-(pr.b + 1)];
        V3 ca, ua;
        core(pr.a, ex, ca, ua);
        const double sl = (SHADOW && !ex) ? slack : 0.0;
        const T3 TB = load12(O.P);
        const double h = pr.bound == 1 ? sa.p[1] : 0.0;
        const V3 e0 = rtmul(TB, axpy(-h, ua, ca) - TB.t), e1 = rtmul(TB, axpy(h, ua, ca) - TB.t);
        const double sx = fmax(fmin(e0.x, e1.x) - O.shape.p[0], -fmax(e0.x, e1.x) - O.shape.p[0]);
        const double sy = fmax(fmin(e0.y, e1.y) - O.shape.p[1], -fmax(e0.y, e1.y) - O.shape.p[1]);
        const double sz = fmax(fmin(e0.z, e1.z) - O.shape.p[2], -fmax(e0.z, e1.z) - O.shape.p[2]);
        const double ra = pr.a >= 0 ? T.geoms[pr.a].bound.br : T.obstacles[-(pr.a + 1)].bound.br;
        const double gap = fmax(sx, fmax(sy, sz)) - ra - pr.margin;
        if (DMIN && gap > 1e-9 + sl) dminLocal = fmin(dminLocal, gap);  // culled = provably apart, not far: the bound is the hint
        return gap > 1e-9 + sl;
      };
      const bool deferSat = mv.ncpairs <= 256;
      unsigned long long surv = 0;
      int nsurv = 0;
      int hit = -1;
      bool ex = exactAll;
      for (int p = 0; p < mv.ncpairs;) {
        const DevPair pr = T.cpairs[p];
        const DevShape& sa = *(const DevShape*)(pr.a >= 0 ? &T.geoms[pr.a].shape : &T.obstacles[-(pr.a + 1)].shape);
        const DevShape& sb = *(const DevShape*)(pr.b >= 0 ? &T.geoms[pr.b].shape : &T.obstacles[-(pr.b + 1)].shape);
        V3 ca, ua, cb, ub;
        core(pr.a, ex, ca, ua);
        core(pr.b, ex, cb, ub);
        if (pr.kind <= PK_CAPSULE_CAPSULE) {
          double d;
          if (pr.kind == PK_SPHERE_SPHERE) { V3 dd = ca - cb; d = sqrt(dot(dd, dd)) - sa.p[0] - sb.p[0]; }
          else if (pr.kind == PK_SPHERE_CAPSULE) d = sqrt(pointSegmentDist2(ca, cb, ub, sb.p[1])) - sa.p[0] - sb.p[0];
          else {
            V3 p1 = axpy(-sa.p[1], ua, ca), q1 = axpy(sa.p[1], ua, ca);
            V3 p2 = axpy(-sb.p[1], ub, cb), q2 = axpy(sb.p[1], ub, cb);
            d = sqrt(segmentSegmentDist2(p1, q1, p2, q2)) - sa.p[0] - sb.p[0];
          }
          const double dm = d - pr.margin;
          if (DMIN) dminLocal = fmin(dminLocal, dm);
          if (SHADOW && !ex) {
            if (dm > HPPB200_COLLISION_THRESHOLD + slack) { ++tests; ++p; continue; }
            if (dm <= HPPB200_COLLISION_THRESHOLD - slack) { ++tests; hit = p; break; }
            ex = true;  // within the slack of the threshold: the same pair again, from the FP64 placements
            continue;
          }
          ++tests;
          if (dm <= HPPB200_COLLISION_THRESHOLD) { hit = p; break; }
          ex = exactAll;
          ++p;
          continue;
        }
        // GJK kinds: conservative cull on the bounding cores (one-sided: the slack only keeps more pairs)
        const double sl = (SHADOW && !ex) ? slack : 0.0, csl = (SHADOW && !ex) ? cullSlack : 0.0;
        double d2;
        if (pr.bound == 1) d2 = pointSegmentDist2(cb, ca, ua, sa.p[1]);
        else { V3 dd = ca - cb; d2 = dot(dd, dd); }
        if (d2 > pr.cull2 + csl) {
          // (DMIN: cores farther apart than the reach, by this much at least; only looked at within 1.1 x the reach: the chains
          //  that matter start millimetres from an obstacle)
          if (DMIN && d2 < 1.21 * pr.cull2) dminLocal = fmin(dminLocal, sqrt(d2) - sqrt(pr.cull2));
          ++p;
          continue;
        }
        if (pr.b < 0 && sb.type == SH_BOX) {
          // The box-face bound runs for the few pairs that pass the first cull, at a different pair per lane: inline it
          // executed at 1.6 of 32 lanes (8 % of the kernel's instructions).  The survivors are noted (up to 8 pair indices of
          // 8 bits) and tested after the loop, every lane walking its own list in step.
          if (deferSat && nsurv < 8) { surv |= (unsigned long long)p << (8 * nsurv); ++nsurv; ++p; continue; }
          if (!boxFaceCull(p, ex)) pushJob(((unsigned long long)i << 32) | (unsigned)p, sjobs, scount, A);
          ++p;
          continue;
        }
        pushJob(((unsigned long long)i << 32) | (unsigned)p, sjobs, scount, A);
        ++p;
      }
      for (int k = 0; k < nsurv; ++k) {
        const int p = (int)((surv >> (8 * k)) & 0xffull);
        if (!boxFaceCull(p, exactAll)) pushJob(((unsigned long long)i << 32) | (unsigned)p, sjobs, scount, A);
      }
      A.valid[i] = hit < 0 ? 1 : 0;
      A.first[i] = (uint32_t)hit;
      A.reason[i] = hit < 0 ? HPPB200_REASON_VALID : HPPB200_REASON_COLLISION;
      if (DMIN) A.dmin[i] = (unsigned long long)__double_as_longlong(fmax(dminLocal, 0.0));  // (gjkJobsKernel runs behind this kernel)
    }
  }
  // ---- flush this block's jobs with ONE global atomic ----
  __syncthreads();
  const int nj = min(*scount, K1_JOBS_PER_BLOCK);
  if (tid == 0 && nj > 0) sbase = atomicAdd(A.njobs, (unsigned long long)nj);
  __syncthreads();
  for (int k = tid; k < nj; k += K1_THREADS)
    if (sbase + k < A.jobs_cap) A.jobs[sbase + k] = sjobs[k];
  const unsigned ts = __reduce_add_sync(0xffffffffu, tests);
  if ((tid & 31) == 0 && stats && ts) atomicAdd(&stats->pair_tests, (unsigned long long)ts);
}

#ifndef K2_THREADS
#define K2_THREADS 128  // 167 registers: 3 CTAs of 128 (12 warps per SM) instead of 1 CTA of 256 (8 warps)
#endif
#ifndef K2_GRID_PER_SM
#define K2_GRID_PER_SM 12
#endif
template <bool DMIN = false>
__global__ void __launch_bounds__(K2_THREADS, 384 / K2_THREADS)
gjkJobsKernel(ModelView mv, ValidateArgs A, DeviceStats* stats) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  unsigned char* sblob = smem + 16;
  stageBlobTMA(sblob, mv.blob, mv.blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);
  unsigned long long total = *A.njobs;
  if (total > A.jobs_cap) total = A.jobs_cap;
  GjkCounters cnt = {0u, 0u, 0u};
  unsigned tests = 0;
  for (unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; j < total;
       j += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long job = A.jobs[j];
    const size_t i = (size_t)(job >> 32);
    const unsigned p = (unsigned)job;
    if (*(volatile uint32_t*)&A.first[i] < p) continue;  // a smaller colliding index is already known
    const DevPair pr = T.cpairs[p];
    const double d = pairLowerBound(T, pr, A.oMg + i, A.n, mv.vertices, cnt);
    ++tests;
    if (d <= HPPB200_COLLISION_THRESHOLD) {
      atomicMin(&A.first[i], p);
      A.valid[i] = 0;
      A.reason[i] = HPPB200_REASON_COLLISION;
    } else if (DMIN) atomicMin(&A.dmin[i], (unsigned long long)__double_as_longlong(d));
  }
  const unsigned ts = __reduce_add_sync(0xffffffffu, tests);
  const unsigned gi = __reduce_add_sync(0xffffffffu, cnt.iterations);
  const unsigned gc = __reduce_add_sync(0xffffffffu, cnt.cap_hits);
  if ((threadIdx.x & 31) == 0 && stats) {
    if (ts) atomicAdd(&stats->pair_tests, (unsigned long long)ts);
    if (gi) atomicAdd(&stats->gjk_iterations, (unsigned long long)gi);
    if (gc) atomicAdd(&stats->gjk_cap_hits, (unsigned long long)gc);
  }
}


// Launch set-up of configPassKernel: the blob view without the continuous-validation tables, the shadow slack inputs, the
// shared-memory size and the kernel instance.  HPPB200_CFG_SHADOW=0 keeps FP64 cores in shared memory (3 CTAs per SM).
namespace {
typedef void (*ConfigPassFn)(ModelView, ValidateArgs, DeviceStats*);
struct ConfigPassSetup {
  ModelView cv;
  uint32_t head, tailOff, tailBytes;
  double hmax, reachMax;
  int forceExact;
  size_t smem;
  ConfigPassFn fn;
};
// firstSample: the pair table is the continuous validation's (the body pairs' geometry pairs in body-pair order,
// SolidSolidCollision's m.pairs flattened), staged with the bounds behind it; the configuration pair table stays out.
ConfigPassSetup configPassSetup(Workspace* ws, bool hostq = false, bool firstSample = false) {
  const ModelView& mv = ws->view;
  const Model* m = ws->model;
  ConfigPassSetup S{};
  const uint32_t tail0 = firstSample ? mv.off_bgpairs : mv.off_lower;
  const uint32_t tail1 = firstSample ? mv.off_cpairs : mv.blob_bytes;
  const uint32_t gap = tail0 - mv.off_bodypairs;
  S.cv = mv;
  S.cv.off_bodypairs = S.cv.off_bgpairs = S.cv.off_coefs = S.cv.off_rcoefs = 0;  // (not staged, not used by the kernel)
  S.cv.off_lower -= gap; S.cv.off_upper -= gap; S.cv.off_weights -= gap; S.cv.off_cpairs -= gap;
  if (firstSample) { S.cv.off_cpairs = mv.off_bgpairs - gap; S.cv.ncpairs = mv.nbgpairs; }
  S.cv.blob_bytes = mv.off_bodypairs + (tail1 - tail0);
  S.head = mv.off_bodypairs;
  S.tailOff = tail0;
  S.tailBytes = tail1 - tail0;
  for (const DevGeom& G : m->geoms)
    if (G.shape.type == HPPB200_SHAPE_CAPSULE) S.hmax = std::max(S.hmax, G.shape.p[1]);
  const DevPair* cp = (const DevPair*)(m->blob.data() + (firstSample ? mv.off_bgpairs : mv.off_cpairs));
  double c2 = 0;
  for (int p = 0; p < S.cv.ncpairs; ++p) c2 = std::max(c2, cp[p].cull2);
  S.reachMax = std::sqrt(c2) * (1.0 + 1e-12);
  bool shadow = true;
  if (const char* v = std::getenv("HPPB200_CFG_SHADOW")) { shadow = std::atoi(v) != 0; S.forceExact = std::atoi(v) == 2; }
  if (hostq && !shadow) throw std::runtime_error("the zero-copy configuration pass exists in the shadow form only");
  S.smem = 16 + S.cv.blob_bytes + (size_t)mv.ngeoms * 6 * K1_THREADS * (shadow ? sizeof(uint32_t) : sizeof(double)) +
           std::max((size_t)K1_JOBS_PER_BLOCK * sizeof(unsigned long long), hostq ? (size_t)K1_THREADS * mv.nq * sizeof(double) : (size_t)0);
  if (S.smem > 227u * 1024u) throw std::runtime_error("model too large for the configuration kernel");
  const bool tree = modelIsTree(m);
  S.fn = hostq ? (tree ? configPassKernel<true, true, true> : configPassKernel<false, true, true>)
         : shadow ? (tree ? configPassKernel<true, true, false> : configPassKernel<false, true, false>)
                  : (tree ? configPassKernel<true, false, false> : configPassKernel<false, false, false>);
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(S.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S.smem));
  return S;
}
void launchConfigPass(const ConfigPassSetup& S, cudaStream_t stream, ValidateArgs A, DeviceStats* stats) {
  A.blob_head = S.head; A.blob_tail_off = S.tailOff; A.blob_tail_bytes = S.tailBytes;
  A.hmax = S.hmax; A.reach_max = S.reachMax; A.force_exact = S.forceExact;
  S.fn<<<(unsigned)((A.n + K1_THREADS - 1) / K1_THREADS), K1_THREADS, S.smem, stream>>>(S.cv, A, stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
}
}  // namespace

// The two kernels on `stream`, scratch given by the caller: oMg [slots*12][n], njobs (one counter), jobs [cap].
// streamJobs / evPass: when given, gjkJobsKernel runs on its own stream behind the event recorded after configPassKernel
// (a latency-bound kernel: it overlaps the next slice's configPassKernel).
static void launchTwoKernelsOn(Workspace* ws, cudaStream_t stream, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid,
                               int32_t* d_first, uint8_t* d_reason, double* oMg, unsigned long long* njobs, unsigned long long* jobs,
                               size_t cap, cudaStream_t streamJobs = nullptr, cudaEvent_t evPass = nullptr, bool hostq = false) {
  const bool tree = modelIsTree(ws->model);
  const ModelView& mv = ws->view;
  ValidateArgs A{};
  A.q = d_q; A.n = n; A.flags = flags; A.valid = d_valid; A.first = (uint32_t*)d_first; A.reason = d_reason;
  A.oMg = oMg;
  A.njobs = njobs;
  A.jobs = jobs;
  A.jobs_cap = cap;
  HPPB200_CUDA_CHECK(cudaMemsetAsync(A.njobs, 0, sizeof(unsigned long long), stream));
  launchConfigPass(configPassSetup(ws, hostq), stream, A, ws->d_stats);
  const size_t smem2 = 16 + mv.blob_bytes;
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(gjkJobsKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
  size_t grid2 = (size_t)ws->num_sms * K2_GRID_PER_SM;
  if (streamJobs) {
    HPPB200_CUDA_CHECK(cudaEventRecord(evPass, stream));
    HPPB200_CUDA_CHECK(cudaStreamWaitEvent(streamJobs, evPass, 0));
    stream = streamJobs;
  }
  gjkJobsKernel<false><<<(unsigned)grid2, K2_THREADS, smem2, stream>>>(mv, A, ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 2;
  ws->stats.configs += n;
}

// First-sample screen of the edge path (kernels_edges.cu::launchValidateEdges): the configurations q(t0) of a batch of
// edges against the CONTINUOUS validation's pair table: first[i] = index (into the body pairs' geometry pairs, reference
// order) of the first colliding pair, which is what BodyPairCollision::computeDistanceLowerBound reports for the first
// sample of an edge (body-pair-collision.cc:175-198: same predicate, same order, nothing validated yet).  In chunks, so
// that the placement scratch stays bounded.  d_njobs [chunks] receives the queue counter of every chunk; the caller
// compares them with firstSampleQueueCap(min(n, chunk)) after synchronising: an overflowed queue voids the screen.
size_t firstSampleChunk() {
  if (const char* v = std::getenv("HPPB200_FS_CHUNK")) return (size_t)std::max(128ll, std::atoll(v));  // (tests: several chunks on a small batch)
  return 262144;
}
size_t firstSampleQueueCap(size_t n) { return 16 * n + 4096; }
// (a model whose tables do not fit the configuration kernel's shared memory is validated without the screen)
bool firstSampleScreenAvailable(Workspace* ws) {
  try {
    (void)configPassSetup(ws, false, true);
    return true;
  } catch (const std::runtime_error&) {
    return false;
  }
}
// Returns whether d_dmin (the scheduling hint, see ValidateArgs::dmin) was filled.
bool launchFirstSampleScreen(Workspace* ws, const double* d_q, size_t n, uint8_t* d_valid, int32_t* d_first, uint8_t* d_reason,
                             unsigned long long* d_njobs /* [chunks] */, unsigned long long* d_dmin /* [n] or null */) {
  const bool tree = modelIsTree(ws->model);
  const ModelView& mv = ws->view;
  ModelView mv2 = mv;
  mv2.off_cpairs = mv.off_bgpairs;
  mv2.ncpairs = mv.nbgpairs;
  ConfigPassSetup S = configPassSetup(ws, false, true);
  if (d_dmin) {  // (the DMIN instances exist in the shadow form; HPPB200_CFG_SHADOW=0: no hint)
    if (S.fn == configPassKernel<true, true, false>) S.fn = configPassKernel<true, true, false, true>;
    else if (S.fn == configPassKernel<false, true, false>) S.fn = configPassKernel<false, true, false, true>;
    else d_dmin = nullptr;
    if (d_dmin) HPPB200_CUDA_CHECK(cudaFuncSetAttribute(S.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S.smem));
  }
  const size_t chunk = std::min(n, firstSampleChunk());
  const size_t slots = (size_t)mv.ngeoms + (tree ? mv.njoints : 0);
  ws->d_scratch.reserve(slots * 12 * chunk);
  const size_t cap = firstSampleQueueCap(chunk);
  ws->d_jobs.reserve(cap + 1);
  const size_t smem2 = 16 + mv.blob_bytes;
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(gjkJobsKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(gjkJobsKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
  const size_t nq = (size_t)ws->model->nq;
  HPPB200_CUDA_CHECK(cudaMemsetAsync(d_njobs, 0, ((n + chunk - 1) / chunk) * sizeof(unsigned long long), ws->stream));
  size_t c = 0;
  for (size_t off = 0; off < n; off += chunk, ++c) {
    const size_t len = std::min(chunk, n - off);
    ValidateArgs A{};
    A.q = d_q + off * nq; A.n = len; A.flags = 0; A.valid = d_valid + off; A.first = (uint32_t*)(d_first + off); A.reason = d_reason + off;
    A.oMg = ws->d_scratch.p;
    A.njobs = d_njobs + c;
    A.jobs = ws->d_jobs.p + 1;
    A.jobs_cap = cap;
    A.dmin = d_dmin ? d_dmin + off : nullptr;
    launchConfigPass(S, ws->stream, A, ws->d_stats);
    if (d_dmin) gjkJobsKernel<true><<<(unsigned)((size_t)ws->num_sms * K2_GRID_PER_SM), K2_THREADS, smem2, ws->stream>>>(mv2, A, ws->d_stats);
    else gjkJobsKernel<false><<<(unsigned)((size_t)ws->num_sms * K2_GRID_PER_SM), K2_THREADS, smem2, ws->stream>>>(mv2, A, ws->d_stats);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    ws->stats.kernel_launches += 2;
  }
  return d_dmin != nullptr;
}

// Jobs that did not fit the queue (pathological scenes): detected on the host, which re-runs the
// batch through the single-kernel path below (still on the device).
static void launchValidateTwoKernel(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid,
                                    int32_t* d_first, uint8_t* d_reason) {
  const bool tree = modelIsTree(ws->model);
  const ModelView& mv = ws->view;
  if (n > 0xffffffffull) throw std::invalid_argument("at most 2^32-1 configurations per call");
  if (!d_first) { ws->d_pair_int.reserve(n); d_first = ws->d_pair_int.p; }
  if (!d_reason) { ws->d_reason_int.reserve(n); d_reason = ws->d_reason_int.p; }
  const size_t slots = (size_t)mv.ngeoms + (tree ? mv.njoints : 0);
  ws->d_scratch.reserve(slots * 12 * n);
  const size_t cap = 2 * n + 4096;
  ws->d_jobs.reserve(cap + 1);
  launchTwoKernelsOn(ws, ws->stream, d_q, n, flags, d_valid, d_first, d_reason, ws->d_scratch.p, ws->d_jobs.p, ws->d_jobs.p + 1, cap);
}

static void launchConfigKernel(Workspace* ws, const double* d_q, size_t n, uint32_t flags, int mode, uint8_t* d_valid,
                               int32_t* d_first, uint8_t* d_reason, double* d_aux) {
  if (n == 0) return;
  const bool tree = modelIsTree(ws->model);
  uint32_t cpb = 64;
  ConfigSmemLayout L = configSmemLayout(ws->view, tree, cpb);
  if (L.total > 227u * 1024u) { cpb = 32; L = configSmemLayout(ws->view, tree, cpb); }
  if (L.total > 227u * 1024u)
    throw std::runtime_error("model too large for the shared-memory configuration kernel (" + std::to_string(L.total) + " B)");
  const unsigned grid = (unsigned)((n + cpb - 1) / cpb);
  auto kern = cpb == 64 ? (tree ? validateConfigsKernel<true, 64> : validateConfigsKernel<false, 64>)
                        : (tree ? validateConfigsKernel<true, 32> : validateConfigsKernel<false, 32>);
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total));
  kern<<<grid, CFG_THREADS, L.total, ws->stream>>>(ws->view, L, d_q, n, flags, mode, d_valid, d_first, d_reason, d_aux,
                                                  ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  if (mode == MODE_VALIDATE) ws->stats.configs += n;
}

// Asynchronous: the queue-overflow counter (more than ~2 surviving GJK pairs per configuration on average) is copied to
// pinned host memory behind the kernels and looked at by finishValidateConfigs AFTER the caller's own synchronisation
// (round 1 synchronised here on every call for a case that does not occur on the bench scenes).
void launchValidateConfigs(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid,
                           int32_t* d_first, uint8_t* d_reason) {
  if (n == 0) return;
  launchValidateTwoKernel(ws, d_q, n, flags, d_valid, d_first, d_reason);
  if (!ws->h_njobs) HPPB200_CUDA_CHECK(cudaHostAlloc((void**)&ws->h_njobs, sizeof(unsigned long long), cudaHostAllocDefault));
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->h_njobs, ws->d_jobs.p, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ws->stream));
}
// After the stream has been synchronised: a batch whose queue overflowed is redone by the single fused kernel, which
// needs no queue (still on the device; synchronises again).
void finishValidateConfigs(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid, int32_t* d_first,
                           uint8_t* d_reason) {
  if (n == 0 || !ws->h_njobs) return;
  ws->last_jobs = *ws->h_njobs;
  if (ws->last_jobs <= 2 * n + 4096) return;
  launchConfigKernel(ws, d_q, n, flags, MODE_VALIDATE, d_valid, d_first, d_reason, nullptr);
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
}

// ------------------------------------------------------------------------------------------------
// hppb200_validate_configs_noalloc: the same two kernels (configPassKernel + gjkJobsKernel) with their scratch
// (job counter, job queue, placement SoA) in a file-scope device array instead of the workspace's cudaMalloc'ed
// buffers, for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
// The array is per device (every GPU that loads the library holds HPPB200_NOALLOC_SCRATCH_BYTES until exit), shared
// by every call and keeps the last call's contents: the job counter is cleared at the start of each call (nothing
// else must start at zero: the queue and the placements are written before they are read), and calls that overlap on
// the GPU race on it.
static __device__ __align__(256) unsigned char g_noallocScratch[HPPB200_NOALLOC_SCRATCH_BYTES];

// The array's name is a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress, once per
// device, with that device current.
static unsigned char* noallocScratchAddress(int device) {
  static std::mutex mu;
  static unsigned char* ptr[64] = {};
  if (device < 0 || device >= 64) throw std::invalid_argument("device index out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[device]) {
    HPPB200_CUDA_CHECK(cudaSetDevice(device));
    void* p = nullptr;
    HPPB200_CUDA_CHECK(cudaGetSymbolAddress(&p, g_noallocScratch));
    if (!p) throw std::runtime_error("cudaGetSymbolAddress returned a null scratch address");
    ptr[device] = (unsigned char*)p;
  }
  return ptr[device];
}
void noallocSetup(int device) { (void)noallocScratchAddress(device); }

void launchValidateConfigsNoAlloc(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid, int32_t* d_first,
                                  uint8_t* d_reason) {
  if (n == 0) return;
  if (!d_q || !d_valid || !d_first || !d_reason) throw std::invalid_argument("hppb200_validate_configs_noalloc: q, valid, first_pair and reason are required");
  const bool tree = modelIsTree(ws->model);
  const ModelView& mv = ws->view;
  const size_t slots = (size_t)mv.ngeoms + (tree ? mv.njoints : 0);
  const size_t cap = 2 * n + 4096;                       // job queue entries, as in launchValidateTwoKernel
  const size_t offJobs = 256, offPlc = (offJobs + cap * sizeof(unsigned long long) + 255) & ~size_t(255);
  const size_t need = offPlc + slots * 12 * n * sizeof(double);
  if (need > (size_t)HPPB200_NOALLOC_SCRATCH_BYTES) {
    const size_t perConfig = slots * 96 + 16;
    throw std::invalid_argument("hppb200_validate_configs_noalloc: " + std::to_string(n) + " configurations need " + std::to_string(need) +
                                " B of scratch, the static array holds " + std::to_string((size_t)HPPB200_NOALLOC_SCRATCH_BYTES) + " B (at most " +
                                std::to_string(((size_t)HPPB200_NOALLOC_SCRATCH_BYTES - 33536) / perConfig) +
                                " configurations per call for this model); use hppb200_validate_configs or split the batch");
  }
  unsigned char* base = noallocScratchAddress(ws->device);
  ValidateArgs A{};
  A.q = d_q; A.n = n; A.flags = flags; A.valid = d_valid; A.first = (uint32_t*)d_first; A.reason = d_reason;
  A.njobs = (unsigned long long*)base;
  A.jobs = (unsigned long long*)(base + offJobs);
  A.jobs_cap = cap;
  A.oMg = (double*)(base + offPlc);
  HPPB200_CUDA_CHECK(cudaMemsetAsync(A.njobs, 0, sizeof(unsigned long long), ws->stream));  // the one thing that must start at zero
  launchConfigPass(configPassSetup(ws), ws->stream, A, ws->d_stats);
  const size_t smem2 = 16 + mv.blob_bytes;
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(gjkJobsKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
  gjkJobsKernel<false><<<(unsigned)((size_t)ws->num_sms * K2_GRID_PER_SM), K2_THREADS, smem2, ws->stream>>>(mv, A, ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 2;
  ws->stats.configs += n;
  // queue overflow: same rule as launchValidateConfigs (the fused single kernel needs no scratch at all)
  unsigned long long njobs = 0;
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&njobs, A.njobs, sizeof(njobs), cudaMemcpyDeviceToHost, ws->stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  if (njobs > cap) launchConfigKernel(ws, d_q, n, flags, MODE_VALIDATE, d_valid, d_first, d_reason, nullptr);
}

// Host-pointer path of hppb200_validate_configs: the batch in slices on three streams (H2D, kernels, D2H), so that the
// H2D copy of slice i + 1 runs under the kernels of slice i and the D2H of slice i - 1 behind them (the monolithic copy -> kernels ->
// copy sequence left the copy engines and the SMs idle in turn: 2.17 ms end to end for 1.14 ms of kernels).
// Every slice has its own placement / queue region; the queue-overflow counters are read once at the end.
// Returns with everything complete (both streams synchronised); ws->ev1 is recorded on ws->stream at the end.
void launchValidateConfigsHost(Workspace* ws, const double* h_q, size_t n, uint32_t flags, uint8_t* h_valid, int32_t* h_first,
                               uint8_t* h_reason) {
  if (n == 0) return;
  if (n > 0xffffffffull) throw std::invalid_argument("at most 2^32-1 configurations per call");
  const bool tree = modelIsTree(ws->model);
  const ModelView& mv = ws->view;
  const size_t nq = (size_t)ws->model->nq;
  const size_t slots = (size_t)mv.ngeoms + (tree ? mv.njoints : 0);
  // slice sizes in whole WAVES of configPassKernel (resident CTAs x 128 configurations; a slice of 1.15 waves costs 2):
  // one wave first (the kernels start after its copy), then doubling up to `maxWaves`; the H2D stream runs ahead of the
  // compute stream (0.86 ns vs 1.09 ns per UR5 configuration), the D2H stream behind it
  // Zero-copy form: h_q is pinned (cudaHostAlloc / cudaHostRegister) and 16-byte aligned, so configPassKernel can pull
  // its rows over PCIe itself (HOSTQ); pageable memory keeps the staged copy.  HPPB200_CFG_ZEROCOPY=1 switches it on.
  bool zeroCopy = false;
  {
    // Off by default: measured 1.38 ms per 1 M UR5 configurations against 1.37 ms for the staged slices (SM-initiated
    // PCIe reads reach ~44 GB/s against ~50 GB/s for the copy engine, and the CTAs of one wave fetch, then compute, together).
    bool want = false, shadow = true;
    if (const char* v = std::getenv("HPPB200_CFG_ZEROCOPY")) want = std::atoi(v) != 0;
    if (const char* v = std::getenv("HPPB200_CFG_SHADOW")) shadow = std::atoi(v) != 0;
    if (want && shadow && ((uintptr_t)h_q & 15u) == 0) {
      cudaPointerAttributes pa{};
      if (cudaPointerGetAttributes(&pa, h_q) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer == (void*)h_q) zeroCopy = true;
      else (void)cudaGetLastError();
    }
  }
  size_t wave;
  {
    const ConfigPassSetup S = configPassSetup(ws, zeroCopy);
    int perSM = 0;
    HPPB200_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, S.fn, K1_THREADS, S.smem));
    wave = (size_t)std::max(1, perSM) * (size_t)ws->num_sms * K1_THREADS;
  }
  // measured (1 M UR5, pinned), one stream for configPassKernel: 4/2 705, 2/2 762, 2/0 766, 1/2 747, 8/2 713 M configurations/s;
  // two alternating streams: 2/2 797, 1/2 845, 4/2 711
  size_t maxWaves = 1, tailWaves = 2;
  if (const char* v = std::getenv("HPPB200_CFG_SLICE_WAVES")) maxWaves = std::max<size_t>(1, (size_t)std::atoll(v));
  if (const char* v = std::getenv("HPPB200_CFG_TAIL_WAVES")) tailWaves = (size_t)std::max(0ll, std::atoll(v));
  // Slices in whole waves: 1, 2, 4, ... (the kernels start after the first, small copy), and `tailWaves` single-wave slices
  // at the end: with the kernels as fast as the copy (0.90 vs 0.97 ms per 1 M UR5 configurations) the call ends one last
  // slice of kernels + D2H after the last byte has arrived, so the last slice is the smallest efficient one.  The
  // fractional wave of the batch rides in the second slice.
  std::vector<size_t> offs, lens;
  {
    const size_t full = n / wave, frac = n - full * wave;
    std::vector<size_t> sz;
    if (full < tailWaves + 2) sz.push_back(n);  // small batch: one slice (below ~4 waves slicing buys nothing)
    else {
      size_t body = full - tailWaves;
      for (size_t w = zeroCopy ? maxWaves : 1; body > 0; w = std::min(maxWaves, w * 2)) {  // (zero-copy: no first copy to wait for)
        const size_t m = std::min(w, body);
        sz.push_back(m * wave);
        body -= m;
      }
      sz[sz.size() > 1 ? 1 : 0] += frac / K1_THREADS * K1_THREADS;  // (slice offsets stay multiples of 128 configurations)
      for (size_t k = 0; k < tailWaves; ++k) sz.push_back(wave);
      sz.back() += frac % K1_THREADS;
    }
    size_t off = 0;
    for (size_t m : sz) { offs.push_back(off); lens.push_back(m); off += m; }
  }
  const size_t ns = offs.size();
  if (!zeroCopy) ws->d_q.reserve(n * nq);
  ws->d_valid.reserve(n);
  ws->d_pair.reserve(n);
  ws->d_reason.reserve(n);
  ws->d_scratch.reserve(slots * 12 * n);
  const size_t capAll = 2 * n + 4096 * ns;
  ws->d_jobs.reserve(capAll + ns + 1);
  if (!ws->stream2) {
    HPPB200_CUDA_CHECK(cudaStreamCreateWithFlags(&ws->stream2, cudaStreamNonBlocking));
    HPPB200_CUDA_CHECK(cudaStreamCreateWithFlags(&ws->stream3, cudaStreamNonBlocking));
    // (a high-priority stream for gjkJobsKernel was measured slower: 655 vs 697 M configurations/s end to end)
    HPPB200_CUDA_CHECK(cudaStreamCreateWithFlags(&ws->stream4, cudaStreamNonBlocking));
    HPPB200_CUDA_CHECK(cudaStreamCreateWithFlags(&ws->stream5, cudaStreamNonBlocking));
    HPPB200_CUDA_CHECK(cudaEventCreateWithFlags(&ws->evJoin, cudaEventDisableTiming));
    HPPB200_CUDA_CHECK(cudaEventCreateWithFlags(&ws->evFork, cudaEventDisableTiming));
  }
  while (ws->evSlice.size() < 3 * ns) {
    cudaEvent_t e;
    HPPB200_CUDA_CHECK(cudaEventCreate(&e));  // (timing kept: HPPB200_CFG_TRACE prints the slice timeline)
    ws->evSlice.push_back(e);
  }
  HPPB200_CUDA_CHECK(cudaEventRecord(ws->evFork, ws->stream));  // (the staging buffers may still be read by earlier work)
  HPPB200_CUDA_CHECK(cudaStreamWaitEvent(ws->stream2, ws->evFork, 0));
  // configPassKernel of consecutive slices on two alternating streams: the next slice's CTAs fill the SMs the last wave of
  // the previous one leaves idle (HPPB200_CFG_K1_STREAMS=1: one stream)
  bool twoK1 = true;
  if (const char* v = std::getenv("HPPB200_CFG_K1_STREAMS")) twoK1 = std::atoi(v) >= 2;
  if (twoK1) HPPB200_CUDA_CHECK(cudaStreamWaitEvent(ws->stream5, ws->evFork, 0));
  unsigned long long* counters = ws->d_jobs.p;       // [ns]
  unsigned long long* queue = ws->d_jobs.p + ns;     // slice k at 2 * off + 4096 * k
  for (size_t k = 0; k < ns; ++k) {
    const size_t off = offs[k], m = lens[k];
    cudaEvent_t evIn = ws->evSlice[3 * k], evDone = ws->evSlice[3 * k + 1], evPass = ws->evSlice[3 * k + 2];
    if (!zeroCopy) HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p + off * nq, h_q + off * nq, m * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream2));
    HPPB200_CUDA_CHECK(cudaEventRecord(evIn, ws->stream2));
    cudaStream_t k1s = (twoK1 && (k & 1)) ? ws->stream5 : ws->stream;
    HPPB200_CUDA_CHECK(cudaStreamWaitEvent(k1s, evIn, 0));
    launchTwoKernelsOn(ws, k1s, zeroCopy ? h_q + off * nq : ws->d_q.p + off * nq, m, flags, ws->d_valid.p + off, ws->d_pair.p + off,
                       ws->d_reason.p + off, ws->d_scratch.p + slots * 12 * off, counters + k, queue + 2 * off + 4096 * k, 2 * m + 4096,
                       ws->stream4, evPass, zeroCopy);
    HPPB200_CUDA_CHECK(cudaEventRecord(evDone, ws->stream4));
    HPPB200_CUDA_CHECK(cudaStreamWaitEvent(ws->stream3, evDone, 0));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(h_valid + off, ws->d_valid.p + off, m, cudaMemcpyDeviceToHost, ws->stream3));
    if (h_first) HPPB200_CUDA_CHECK(cudaMemcpyAsync(h_first + off, ws->d_pair.p + off, m * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream3));
    if (h_reason) HPPB200_CUDA_CHECK(cudaMemcpyAsync(h_reason + off, ws->d_reason.p + off, m, cudaMemcpyDeviceToHost, ws->stream3));
  }
  HPPB200_CUDA_CHECK(cudaEventRecord(ws->evJoin, ws->stream3));
  HPPB200_CUDA_CHECK(cudaStreamWaitEvent(ws->stream, ws->evJoin, 0));
  std::vector<unsigned long long> h_cnt(ns);
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(h_cnt.data(), counters, ns * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ws->stream));
  HPPB200_CUDA_CHECK(cudaEventRecord(ws->ev1, ws->stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  if (std::getenv("HPPB200_CFG_TRACE")) {
    for (size_t k = 0; k < ns; ++k) {
      float a = 0, b = 0;
      float c = 0;
      cudaEventElapsedTime(&a, ws->ev0, ws->evSlice[3 * k]);
      cudaEventElapsedTime(&b, ws->ev0, ws->evSlice[3 * k + 1]);
      cudaEventElapsedTime(&c, ws->ev0, ws->evSlice[3 * k + 2]);
      std::fprintf(stderr, "[cfg] slice %zu: %zu configurations, H2D done at %.3f ms, configPass at %.3f, gjkJobs at %.3f ms\n", k, lens[k], a, c, b);
    }
    float t = 0;
    cudaEventElapsedTime(&t, ws->ev0, ws->ev1);
    std::fprintf(stderr, "[cfg] all copied back at %.3f ms\n", t);
  }
  ws->last_jobs = 0;
  for (size_t k = 0; k < ns; ++k) {
    const size_t off = offs[k], m = lens[k];
    ws->last_jobs += h_cnt[k];
    if (h_cnt[k] <= 2 * m + 4096) continue;
    // queue overflow in this slice: the fused single kernel needs no queue (same rule as launchValidateConfigs)
    if (zeroCopy) {
      ws->d_q.reserve(n * nq);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p + off * nq, h_q + off * nq, m * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    }
    launchConfigKernel(ws, ws->d_q.p + off * nq, m, flags, MODE_VALIDATE, ws->d_valid.p + off, ws->d_pair.p + off, ws->d_reason.p + off, nullptr);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(h_valid + off, ws->d_valid.p + off, m, cudaMemcpyDeviceToHost, ws->stream));
    if (h_first) HPPB200_CUDA_CHECK(cudaMemcpyAsync(h_first + off, ws->d_pair.p + off, m * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    if (h_reason) HPPB200_CUDA_CHECK(cudaMemcpyAsync(h_reason + off, ws->d_reason.p + off, m, cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaEventRecord(ws->ev1, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  }
}

// Few configurations (one-item calls): the CTA-cooperative kernel, whose warps share out the pairs of a configuration.
void launchValidateConfigsSmall(Workspace* ws, const double* d_q, size_t n, uint32_t flags, uint8_t* d_valid, int32_t* d_first, uint8_t* d_reason) {
  launchConfigKernel(ws, d_q, n, flags, MODE_VALIDATE, d_valid, d_first, d_reason, nullptr);
}
void launchGeometryPlacements(Workspace* ws, const double* d_q, size_t n, double* d_oMg) {
  launchConfigKernel(ws, d_q, n, 0, MODE_PLACEMENTS, nullptr, nullptr, nullptr, d_oMg);
}
void launchPairDistances(Workspace* ws, const double* d_q, size_t n, double* d_dist) {
  launchConfigKernel(ws, d_q, n, 0, MODE_DISTANCES, nullptr, nullptr, nullptr, d_dist);
}


// ------------------------------------------------------------------------------------------------
// CollisionValidation with computeAllContacts (reference src/collision-validation.cc:72-89): the report lists the first
// colliding pair and every LATER colliding pair of the same table (cPairs_, or pPairs_ when the first hit is there).
// One thread per configuration over the distances of MODE_DISTANCES (distance - margin of every pair in table order).
__global__ void allCollisionsKernel(const double* __restrict__ dist, size_t n, int npairs, int nparam, uint8_t* valid, int32_t* first,
                                    uint32_t* mask, int32_t* count) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int words = (npairs + 31) / 32, nc = npairs - nparam;
  const double* d = dist + i * (size_t)npairs;
  int f = -1;
  for (int p = 0; p < npairs && f < 0; ++p)
    if (d[p] <= HPPB200_COLLISION_THRESHOLD) f = p;
  uint32_t* m = mask + i * (size_t)words;
  for (int w = 0; w < words; ++w) m[w] = 0u;
  int c = 0;
  if (f >= 0) {
    const int end = f < nc ? nc : npairs;  // the table the first hit is in
    for (int p = f; p < end; ++p)
      if (d[p] <= HPPB200_COLLISION_THRESHOLD) { m[p >> 5] |= 1u << (p & 31); ++c; }
  }
  valid[i] = f < 0 ? 1 : 0;
  if (first) first[i] = f;
  if (count) count[i] = c;
}

void launchAllCollisions(Workspace* ws, const double* d_q, size_t n, uint8_t* d_valid, int32_t* d_first, uint32_t* d_mask, int32_t* d_count) {
  if (n == 0) return;
  const int np = (int)ws->model->cPairs.size();
  ws->d_scratch.reserve((size_t)np * n);
  launchConfigKernel(ws, d_q, n, 0, MODE_DISTANCES, nullptr, nullptr, nullptr, ws->d_scratch.p);
  allCollisionsKernel<<<(unsigned)((n + 127) / 128), 128, 0, ws->stream>>>(ws->d_scratch.p, n, np, ws->model->nParameterized, d_valid, d_first,
                                                                          d_mask, d_count);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
}

// DistanceBetweenObjects::computeDistances (reference src/distance-between-objects.cc:62-77): the distance (no security
// margin) of every (inner object of a body, obstacle) pair, obstacle-major in addObstacle order (:41-53).
// Thread per (configuration, pair) over the placements of MODE_PLACEMENTS.
__global__ void distancePairsKernel(ModelView mv, const DevPair* __restrict__ pairs, int npairs, const double* __restrict__ oMg, size_t n,
                                    double* __restrict__ out) {
  const Tables T = makeTables(mv.blob, mv);
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * (size_t)npairs) return;
  const size_t i = idx / (size_t)npairs;
  const int p = (int)(idx - i * (size_t)npairs);
  GjkCounters cnt = {0u, 0u, 0u};
  out[idx] = pairLowerBound(T, pairs[p], oMg + i * (size_t)mv.ngeoms * 12, 1, mv.vertices, cnt);
}

void launchDistanceBetweenObjects(Workspace* ws, const double* d_q, size_t n, double* d_out) {
  const std::vector<DevPair>& pairs = ws->model->dboPairs;
  if (n == 0 || pairs.empty()) return;
  if (!ws->dboUploaded) {
    ws->d_dbo.reserve(pairs.size());
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_dbo.p, pairs.data(), pairs.size() * sizeof(DevPair), cudaMemcpyHostToDevice, ws->stream));
    ws->dboUploaded = true;
  }
  ws->d_scratch.reserve((size_t)ws->view.ngeoms * 12 * n);
  launchConfigKernel(ws, d_q, n, 0, MODE_PLACEMENTS, nullptr, nullptr, nullptr, ws->d_scratch.p);
  const size_t total = n * pairs.size();
  distancePairsKernel<<<(unsigned)((total + 127) / 128), 128, 0, ws->stream>>>(ws->view, ws->d_dbo.p, (int)pairs.size(), ws->d_scratch.p, n, d_out);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
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
