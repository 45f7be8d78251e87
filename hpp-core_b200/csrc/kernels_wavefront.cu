// Wavefront engine for continuousValidation::Progressive over many StraightPath edges
// (reference src/continuous-validation/progressive.cc:78-129): the production path for large batches.
//
// The sequential algorithm per edge is: sample t -> FK -> walk the body pairs in the current order,
// skipping those whose validated interval already covers the running interval, evaluating the others
// (narrow phase + velocity bound => validated interval) -> t = end of the intersection -> next sample.
// One sample of one edge is little work, samples of one edge are strictly ordered, but edges are
// independent.  So the engine advances ALL resident edges by (at most) one sample per ROUND, and every
// round is two uniform kernels with their own register budget and occupancy:
//
//   wfControlKernel  one thread per edge slot.  REPLAY: follows validateIntervals
//       (include/hpp/core/continuous-validation.hh:267-298) and BodyPairCollision::validateConfiguration
//       (src/continuous-validation/body-pair-collision.cc:45-99) over the pair positions, consuming the
//       distances computed last round.  If every pair the reference evaluates had its distance, the sample
//       is complete: smallest-interval-first swap (src/continuous-validation.cc:144-146), t advances
//       (progressive.cc:104-117), the edge finishes (slot refilled from the global edge counter, ballot +
//       prefix sums) or FK runs for the next sample.  EMIT: the pair the reference evaluates next plus the
//       predicted candidates after it are appended to the job lists (one job per GEOMETRY pair).
//   wfNarrowKernel   one thread per job.  GJK jobs run one iteration per loop trip and finished lanes are
//       refilled from the job list (iteration-level refill keeps warps full although iteration counts
//       differ); closed-form jobs are one thread each.  Result = distance_lower_bound of the pair, 8 bytes.
//
// Which tests to queue is a PREDICTION; what the results mean is decided by the replay, which follows the
// reference order exactly: a pair the reference skips is skipped even if its distance was
// computed, a pair it evaluates whose distance is missing stops the replay until next round.  The distance
// of (edge, t, pair) is a pure function, so flags, parameters and intervals are bit-identical to the
// sequential algorithm whatever was predicted.
//
// State lives in global memory, position-major SoA over the slots (coalesced for the thread-per-slot scan):
// the validated interval and pair id are stored BY ORDER POSITION and swapped together, so the scan reads
// lohi[p][slot] without an indirection.  valid_ (whole path validated) is encoded as lo = -inf.
// Progressive keeps each pair's validated set connected, so one closed interval per pair;
// an insert that would open a second interval hands the edge to the walker (never silently accepted).
// Placements: AoS per slot (96 contiguous bytes per geometry, read by the jobs).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "dev_intervals.cuh"

namespace hppb200 {

constexpr int WF_RB = 12;       // body pairs emitted per slot per round (UR5-class models; WFArgs::rb is the run's value)
constexpr int WF_RG = 16;       // geometry-pair jobs (= result slots) per slot per round (WFArgs::rg)
constexpr int WF_RB_BIG = 32;   // models with more than 64 body pairs: the first samples of an edge evaluate most pairs
constexpr int WF_RG_BIG = 48;
constexpr int WF_CTHREADS = 128;
constexpr int WF_NTHREADS = 128;
#ifndef WF_CHUNK_JOBS
#define WF_CHUNK_JOBS 32  // (128: 68.5 ms of narrow phase per 1 M UR5 edges, 64: 62.7, 32: 61.4: the last chunks of a round are the imbalance)
#endif
constexpr int WF_CHUNK = WF_CHUNK_JOBS;   // jobs a warp takes from the list per atomic

struct WFRoundCtl {
  unsigned nG, nC, headG, headC;
};

struct WFArgs {
  const double* q1;
  const double* q2;
  const double* L;
  size_t n;
  int reverse;
  double tol;
  uint32_t max_iter;
  uint8_t* valid;
  double* t_fail;
  double* t_last;
  int32_t* pair;
  uint8_t* reason;
  int overflow_reason;
  double alpha;
  unsigned long long* counter;  // next edge to hand out
  const uint32_t* perm;         // hand-out order (longest edge first, kernels_sort.cu); null: input order
  int handover;                 // Progressive: slots that reach a sample boundary leave for the CTA-per-edge kernel
  uint32_t* hoList;             // ... their slot numbers
  unsigned* hoCount;
  uint32_t E;                   // edge slots (allocation stride of the [..][E] arrays)
  uint32_t Ecur;                // slots the grids cover; shrinks when the live slots are compacted
  int fkSlots;
  double2* lohi;       // [NP][E] by order position
  uint16_t* ord;       // [E][ordStride] pair id at order position, one record per slot (its reads stay in L1)
  uint32_t ordStride;  // NP rounded up to 16 entries (32 bytes)
  double* vmax;        // [NP][E] by pair id: SolidSolidCollision::computeMaximalVelocity, fixed per edge
  double2* vht;        // Progressive: [NP][E] by pair id: (Vmax, 2 tol / Vmax), both fixed per edge (body-pair-collision.cc:80-84)
  double* sv;          // [nv][E] per-joint velocity norms (refill scratch)
  double* st_t;        // [E] ...
  double* st_last;
  double* st_hp;
  double* st_ilo;
  double* st_ihi;
  double* st_L;
  uint32_t* st_edge;
  uint32_t* st_iters;
  uint32_t* st_pos;    // pos | smallest << 16
  uint8_t* st_flags;   // bit0 active, bit1 retired, bits 2-3 `finished`
  uint8_t* st_nEm;
  unsigned long long* st_vmask;  // [W][E] valid_ bits by order position
  unsigned long long* st_cbits;  // Dichotomy: [4 * W][E] bit planes of the pairs' list lengths
  double2* ivl;        // Dichotomy: [NP * WD_K][E] validated intervals by order position
  double2* sub;        // Dichotomy: [WD_KS][E] valid subset of the edge
  uint8_t* st_subCnt;  // Dichotomy: [E]
  // control form 3 (Progressive): the emission record and the results slot-major, so that one thread pulls its slot's
  // rows into shared memory with a few 16-byte cp.async at kernel start (one latency instead of one per use)
  uint32_t* emA;       // [E][rb]: position | first result << 8 | pair id << 16 | geometry pairs << 24   (null: emPos / emRes below)
  double* resA;        // [E][rg] (null: res below)
  uint16_t* emPos;     // [WF_RB][E] order position of the k-th emitted body pair
  uint8_t* emRes;      // [WF_RB][E] its first result slot
  double2* fold;       // [WF_RB][E] its folded result: validated interval (a, b), or (geometry pair, -inf)
  double* res;         // [WF_RG][E]
  double* plc;         // [E][fkSlots*12]
  unsigned long long* jobsG;  // [E*WF_RG]
  unsigned long long* jobsC;
  WFRoundCtl* ctl;     // [2], by round parity
  unsigned* alive;     // [rounds per batch]
  int round;           // global round number
  int aliveIdx;
  int rb, rg;          // per-round caps: body pairs / geometry-pair jobs per slot
  int refill_min;      // idle lanes in a warp before new GJK jobs are fetched (32 = whole-warp batches)
};

// ------------------------------------------------------------------------------------------------
// Control kernel: one thread per edge slot.  ONE scan over the pair positions per round:
//   REPLAY mode  folds the distances computed last round exactly like validateIntervals and, position by
//                position, marks the candidates of the NEXT sample (predicted from the running intersection);
//   EMIT mode    (after the first pair the reference evaluates whose distance is missing) marks the
//                candidates of the CURRENT sample from that pair on.
// The interval stream lohi[p][slot] is moved global -> shared with cp.async, two chunks of WF_CH positions
// in flight per thread, so a scan exposes NP / WF_CH memory latencies instead of NP.
// shared memory: [0,8) mbarrier | [16,48) block counters | chunk ring | model blob prefix
constexpr int WF_CH = 4;
constexpr uint32_t WF_RING_BYTES = 2u * WF_CH * WF_CTHREADS * 16u;

__device__ __forceinline__ void cpAsync16(void* sdst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(sdst)), "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void cpAsyncCommit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cpAsyncWait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// position bit masks, W 64-bit words (W = 1 covers UR5-class models: one shift per update)
typedef unsigned long long u64;
template <int W>
__device__ __forceinline__ bool maskGet(const u64 (&m)[W], int p) {
  if (W == 1) return (m[0] >> p) & 1ull;
  bool r = false;
#pragma unroll
  for (int w = 0; w < W; ++w)
    if ((p >> 6) == w) r = (m[w] >> (p & 63)) & 1ull;
  return r;
}
template <int W>
__device__ __forceinline__ void maskSet(u64 (&m)[W], int p, bool v) {  // OR the bit in
  if (W == 1) { m[0] |= (u64)(v ? 1 : 0) << p; return; }
#pragma unroll
  for (int w = 0; w < W; ++w)
    if ((p >> 6) == w) m[w] |= (u64)(v ? 1 : 0) << (p & 63);
}
template <int W>
__device__ __forceinline__ void maskPut(u64 (&m)[W], int p, bool v) {  // overwrite the bit
  if (W == 1) { m[0] = (m[0] & ~(1ull << p)) | ((u64)(v ? 1 : 0) << p); return; }
#pragma unroll
  for (int w = 0; w < W; ++w)
    if ((p >> 6) == w) m[w] = (m[w] & ~(1ull << (p & 63))) | ((u64)(v ? 1 : 0) << (p & 63));
}

// ---- pieces shared by the Progressive and Dichotomy control kernels ---------------------------------------
// Per-edge tables of the edges that just entered a slot, the whole warp working on one new edge at a time
// (lane = joint, then lane = body pair): intervals empty, order = construction order, velocity norms, maximal
// velocities.  Called by all lanes; `got` marks the lanes whose slot received an edge.
__device__ __forceinline__ void wfNewEdgeTables(const ModelView& mv, const WFArgs& A, const Tables& T, uint32_t e, int lane, bool got,
                                                uint32_t edge, double L) {
  const unsigned FULL = 0xffffffffu;
  const uint32_t E = A.E;
  const int NP = mv.nbodypairs;
  const double INF = CUDART_INF;
  unsigned gm = __ballot_sync(FULL, got);
  while (gm) {
    const int srcLane = __ffs(gm) - 1;
    gm &= gm - 1u;
    const uint32_t es = e - (uint32_t)lane + (uint32_t)srcLane;
    const uint32_t ed = __shfl_sync(FULL, edge, srcLane);
    const double Ls = __shfl_sync(FULL, L, srcLane);
    const double* q1 = A.q1 + (size_t)ed * mv.nq;
    const double* q2 = A.q2 + (size_t)ed * mv.nq;
    if (Ls != 0) {
      // per-joint norms of Vb = |q2 - q1| / L (StraightPath::impl_velocityBound, straight-path.cc:132-140)
      for (int j = lane; j < mv.njoints; j += 32) {
        const DevJoint& J = T.joints[j];
        const int nvj = J.type == HPPB200_JOINT_FREEFLYER ? 6 : 1;
        A.sv[(size_t)J.idx_v * E + es] = jointVelocityNorm(q1, q2, J.idx_q, nvj, Ls);
      }
    }
    __syncwarp();
    // SolidSolidCollision::computeMaximalVelocity (solid-solid-collision.cc:76-99), fixed per edge and pair
    for (int id = lane; id < NP; id += 32) {
      const DevBodyPair& B = T.bodypairs[id];
      double Vm;
      if (Ls == 0) Vm = INF;  // BodyPairCollision::setupPath, zero-length path
      else {
        double ab = 0;
        for (int c = B.coef_begin; c < B.coef_end; ++c) ab += T.coefs[c].value * __ldcg(A.sv + (size_t)T.coefs[c].idx_v * E + es);
        Vm = ab;
        if (B.joint_b != 0) {
          double ba = 0;
          for (int c = B.rcoef_begin; c < B.rcoef_end; ++c) ba += T.rcoefs[c].value * __ldcg(A.sv + (size_t)T.rcoefs[c].idx_v * E + es);
          if (ba < ab) Vm = ba;
        }
      }
      if (A.vht) A.vht[(size_t)id * E + es] = make_double2(Vm, (Vm != 0) ? 2 * A.tol / Vm : INF);
      else A.vmax[(size_t)id * E + es] = Vm;
      if (A.lohi) A.lohi[(size_t)id * E + es] = make_double2(INF, -INF);
      A.ord[(size_t)es * A.ordStride + id] = (uint16_t)id;
    }
  }
  __syncwarp();
}

// One job per geometry pair of every candidate body pair, in position order; job list slots by warp prefix
// sums, one shared-memory atomic per warp, one global atomic per CTA and list.  Called by all threads of the CTA.
template <int W>
__device__ __forceinline__ void wfEmitJobs(const WFArgs& A, const Tables& T, uint32_t e, int lane, bool active, const u64 (&cand)[W],
                                           int& nEm, unsigned* sCnt, WFRoundCtl* ctl) {
  const unsigned FULL = 0xffffffffu;
  const uint32_t E = A.E;
  uint32_t jl[WF_RG_BIG];  // geometry pair | GJK flag << 31, in result order
  int nr = 0, nGl = 0, nCl = 0;
  if (active) {
    int nb = 0;
    bool full = false;
#pragma unroll
    for (int w = 0; w < W; ++w) {
      u64 mw = cand[w];
      while (mw && !full) {
        const int p = w * 64 + __ffsll((long long)mw) - 1;
        mw &= mw - 1ull;
        const int id = A.ord[(size_t)e * A.ordStride + p];
        const DevBodyPair& B = T.bodypairs[id];
        const int ng = B.pair_end - B.pair_begin;
        if (nb >= A.rb || nr + ng > A.rg) { full = true; break; }
        if (A.emA) A.emA[(size_t)e * A.rb + nb] = (uint32_t)p | ((uint32_t)nr << 8) | ((uint32_t)id << 16) | ((uint32_t)ng << 24);
        else {
          A.emPos[(size_t)nb * E + e] = (uint16_t)p;
          A.emRes[(size_t)nb * E + e] = (uint8_t)nr;
        }
        for (int gp = B.pair_begin; gp < B.pair_end; ++gp) {
          const bool g = T.bgpairs[gp].kind == PK_GJK;
          jl[nr++] = (uint32_t)gp | (g ? 0x80000000u : 0u);
          if (g) ++nGl; else ++nCl;
        }
        ++nb;
      }
    }
    nEm = nb;
  }
  int gOff = nGl, cOff = nCl;  // inclusive scans
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int g2 = __shfl_up_sync(FULL, gOff, o), c2 = __shfl_up_sync(FULL, cOff, o);
    if (lane >= o) { gOff += g2; cOff += c2; }
  }
  // one global atomic per warp and list (a CTA-wide prefix cost two barriers that every warp of the CTA waited at:
  // 12 % of the control kernel's stall samples)
  unsigned wG = 0, wC = 0;
  if (lane == 31) {
    wG = gOff ? atomicAdd(&ctl->nG, (unsigned)gOff) : 0u;
    wC = cOff ? atomicAdd(&ctl->nC, (unsigned)cOff) : 0u;
  }
  wG = __shfl_sync(FULL, wG, 31);
  wC = __shfl_sync(FULL, wC, 31);
  size_t gi = (size_t)wG + (unsigned)(gOff - nGl);
  size_t ci = (size_t)wC + (unsigned)(cOff - nCl);
  for (int i = 0; i < nr; ++i) {
    const uint32_t j = jl[i];
    const unsigned long long job =
        (unsigned long long)e | ((unsigned long long)(j & 0x00ffffffu) << 32) | ((unsigned long long)i << 56);
    if (j & 0x80000000u) A.jobsG[gi++] = job;
    else A.jobsC[ci++] = job;
  }
}

// alive counter of the round + statistics: warp reduce, shared-memory atomics, one global atomic per CTA
__device__ __forceinline__ void wfPublish(const WFArgs& A, DeviceStats* stats, unsigned* sCnt, int lane, bool retired, unsigned samples,
                                          unsigned discarded, unsigned caps, unsigned overflows) {
  const unsigned FULL = 0xffffffffu;
  const int nAlive = __syncthreads_count(!retired);  // a free slot retires only after a failed refill
  const unsigned sm = __reduce_add_sync(FULL, samples), dc = __reduce_add_sync(FULL, discarded);
  const unsigned cp = __reduce_add_sync(FULL, caps), ov = __reduce_add_sync(FULL, overflows);
  if (lane == 0) {
    if (sm) atomicAdd(&sCnt[4], sm);
    if (dc) atomicAdd(&sCnt[5], dc);
    if (cp) atomicAdd(&sCnt[6], cp);
    if (ov) atomicAdd(&sCnt[7], ov);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (nAlive) atomicAdd(A.alive + A.aliveIdx, (unsigned)nAlive);
    if (stats) {
      if (sCnt[4]) atomicAdd(&stats->edge_samples, (unsigned long long)sCnt[4]);
      if (sCnt[5]) atomicAdd(&stats->discarded_tests, (unsigned long long)sCnt[5]);
      if (sCnt[6]) atomicAdd(&stats->iter_cap_hits, (unsigned long long)sCnt[6]);
      if (sCnt[7]) atomicAdd(&stats->overflow, (unsigned long long)sCnt[7]);
      if (blockIdx.x == 0) atomicAdd(&stats->engine_ticks, 1ull);
    }
  }
}

// bits c0 .. c0 + WF_CH - 1 of a position mask, and OR-ing such a chunk back (one shift per chunk for W = 1)
template <int W>
__device__ __forceinline__ unsigned chunkBits(const u64 (&m)[W], int c0) {
  if (W == 1) return c0 < 64 ? (unsigned)(m[0] >> c0) & ((1u << WF_CH) - 1u) : 0u;
  unsigned r = 0;
#pragma unroll
  for (int k = 0; k < WF_CH; ++k) r |= (maskGet<W>(m, c0 + k) ? 1u : 0u) << k;
  return r;
}
template <int W, int N>
__device__ __forceinline__ unsigned chunkBitsN(const u64 (&m)[W], int c0) {  // bits c0 .. c0 + N - 1
  if (W == 1) return c0 < 64 ? (unsigned)(m[0] >> c0) & ((1u << N) - 1u) : 0u;
  unsigned r = 0;
#pragma unroll
  for (int k = 0; k < N; ++k) r |= (maskGet<W>(m, c0 + k) ? 1u : 0u) << k;
  return r;
}
template <int W>
__device__ __forceinline__ void chunkOr(u64 (&m)[W], int c0, unsigned bits) {
  if (W == 1) { m[0] |= (u64)bits << c0; return; }
#pragma unroll
  for (int k = 0; k < WF_CH; ++k) maskSet<W>(m, c0 + k, (bits >> k) & 1u);
}

template <int W>
__global__ void __launch_bounds__(WF_CTHREADS)
wfControlKernel(ModelView mv, WFArgs A, DeviceStats* stats) {
  extern __shared__ __align__(16) unsigned char smem[];  // @region setup
  uint64_t* bar = (uint64_t*)smem;
  unsigned* sCnt = (unsigned*)(smem + 16);  // 0 nG, 1 nC, 2 baseG, 3 baseC, 4 samples, 5 discarded, 6 caps, 7 overflows
  double2* ring = (double2*)(smem + 48) + threadIdx.x;  // entry (buffer b, k) at ring[(b * WF_CH + k) * WF_CTHREADS]
  unsigned char* sblob = smem + 48 + WF_RING_BYTES;
  if (threadIdx.x < 8) sCnt[threadIdx.x] = 0u;
  stageBlobTMA(sblob, mv.blob, mv.edge_blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);

  const unsigned FULL = 0xffffffffu;
  const uint32_t E = A.E;
  const uint32_t e = blockIdx.x * WF_CTHREADS + threadIdx.x;  // E is a multiple of the CTA size
  const int lane = threadIdx.x & 31;
  const int NP = mv.nbodypairs;
  const double INF = CUDART_INF;
  const bool reverse = A.reverse != 0;
  WFRoundCtl* ctl = A.ctl + (A.round & 1);
  if (e == 0) {  // the other parity was consumed by the previous round's narrow phase
    WFRoundCtl* nx = A.ctl + ((A.round + 1) & 1);
    nx->nG = 0; nx->nC = 0; nx->headG = 0; nx->headC = 0;
  }

  const uint8_t flags0 = A.st_flags[e];
  bool active = (flags0 & 1) != 0, retired = (flags0 & 2) != 0;
  int finished = (flags0 >> 2) & 3;
  double t = 0, lastValid = 0, hp = INF, ilo = -INF, ihi = INF, L = 0;
  uint32_t edge = 0, iters = 0;
  int pos = 0, smallest = 0, nEm = 0;
  if (active) {
    t = A.st_t[e]; lastValid = A.st_last[e]; hp = A.st_hp[e]; ilo = A.st_ilo[e]; ihi = A.st_ihi[e]; L = A.st_L[e];
    edge = A.st_edge[e]; iters = A.st_iters[e];
    const uint32_t ps = A.st_pos[e];
    pos = (int)(ps & 0xffffu); smallest = (int)(ps >> 16);
    nEm = A.st_nEm[e];
  }
  unsigned samples = 0, caps = 0, overflows = 0, discarded = 0;
  bool needSample = false, needFK = false;
  u64 cand[W], vmask[W];  // by order position: candidate for the next narrow phase / valid_
#pragma unroll
  for (int w = 0; w < W; ++w) { cand[w] = 0ull; vmask[w] = active ? A.st_vmask[(size_t)w * E + e] : 0ull; }

  auto finishEdge = [&](bool ok, int rs, double tf, double tl, int pr) {
    A.valid[edge] = ok ? 1 : 0;
    if (A.t_fail) A.t_fail[edge] = tf;
    if (A.t_last) A.t_last[edge] = tl;
    if (A.pair) A.pair[edge] = pr;
    if (A.reason) A.reason[edge] = (uint8_t)rs;
    active = false;
  };

  // ---- the scan ---------------------------------------------------------------------------------  @region scan
  if (active) {
    int cursor = 0;
    int nextP = nEm > 0 ? (int)A.emPos[e] : -1;
    bool emit = false, stop = false;
    double ql = 0, qh = 0;
    const double2* src = A.lohi + e;
    auto issue = [&](int c0, int b) {  // chunk starting at position c0 into ring buffer b (valid_ positions need no data)
      const unsigned vch = chunkBits<W>(vmask, c0);
      const double2* g = src + (size_t)c0 * E;
#pragma unroll
      for (int k = 0; k < WF_CH; ++k, g += E)
        if (c0 + k < NP && !((vch >> k) & 1u)) cpAsync16(ring + (b * WF_CH + k) * WF_CTHREADS, g);
      cpAsyncCommit();
    };
    issue(pos, 0);
    int b = 0;
    for (int c0 = pos; c0 < NP && !stop; c0 += WF_CH, b ^= 1) {
      issue(c0 + WF_CH, b ^ 1);  // (an empty group past the end)
      cpAsyncWait<1>();
      const unsigned vch = chunkBits<W>(vmask, c0);
      unsigned cbits = 0u, vbits = 0u;  // candidates / newly valid_ pairs of this chunk
#pragma unroll
      for (int k = 0; k < WF_CH; ++k) {
        const int p = c0 + k;
        if (p >= NP || stop) break;
        if ((vch >> k) & 1u) {
          // valid_ on the whole path: interval = timeRange (body-pair-collision.cc:51-55).  Never a candidate,
          // so never has a queued result.
          if (!emit) {
            if (ihi > L) smallest = p;
            ilo = (ilo < 0.0) ? 0.0 : ilo;
            ihi = (L < ihi) ? L : ihi;
          }
          continue;
        }
        const double2 lh = ring[(b * WF_CH + k) * WF_CTHREADS];
        if (emit) {
          cbits |= (!(lh.x <= ql && qh <= lh.y) ? 1u : 0u) << k;
          continue;
        }
        const bool have = nextP == p;
        const int kk = cursor;
        if (have) {
          ++cursor;
          nextP = cursor < nEm ? (int)A.emPos[(size_t)cursor * E + e] : -1;
        }
        double nl = lh.x, nh = lh.y;  // this pair's validated interval after this position
        if (lh.x <= ilo && ihi <= lh.y) {  // validInterval_ contains the running interval (:56-63)
          if (have) ++discarded;
        } else if (have) {
          const double2 rr = A.fold[(size_t)kk * E + e];
          const double a = rr.x, bb = rr.y;
          if (bb == -INF) {  // collision: a carries the geometry pair
            finishEdge(false, HPPB200_REASON_COLLISION, t, lastValid, (int)a);
            stop = true;
            continue;
          }
          const double l = lh.x, h = lh.y;
          if (!(l > h) && (h < a || bb < l)) {
            // validInterval_.insert (:90) would open a second interval: hand the edge to the walker
            ++overflows;
            finishEdge(false, A.overflow_reason, t, lastValid, -1);
            stop = true;
            continue;
          }
          nl = a; nh = bb;
          if (!(l > h)) { nl = (a < l) ? a : l; nh = (h > bb) ? h : bb; }
          if (nl <= 0.0 && L <= nh) { nl = -INF; nh = INF; vbits |= 1u << k; }  // valid_ (:92-96), lo = -inf
          A.lohi[(size_t)p * E + e] = make_double2(nl, nh);
          if (ihi > bb) smallest = p;  // validateIntervals: fold into the running intersection
          ilo = (ilo < a) ? a : ilo;
          ihi = (bb < ihi) ? bb : ihi;
        } else {
          // the reference evaluates this pair now and its distance is not there yet
          emit = true;
          pos = p;
          ql = fmax(ilo, t - hp);
          qh = fmin(ihi, t + hp);
#pragma unroll
          for (int w = 0; w < W; ++w) cand[w] = 0ull;
          cbits = 1u << k;
          continue;
        }
        // candidate of the NEXT sample: predicted t' = end of the running intersection, window alpha x step
        const double tn = reverse ? ilo : ihi;
        const double hn = A.alpha * fabs(tn - t);
        cbits |= (!(nl <= tn - hn && tn + hn <= nh) ? 1u : 0u) << k;
      }
      chunkOr<W>(cand, c0, cbits);
      if (vbits) chunkOr<W>(vmask, c0, vbits);
    }
    cpAsyncWait<0>();
    if (active && !emit) {
      // sample complete.  src/continuous-validation.cc:144-146: smallest interval first
      if (NP > 1 && smallest != 0) {
        const size_t i0 = e, is = (size_t)smallest * E + e;
        const double2 x0 = A.lohi[i0], xs = A.lohi[is];
        uint16_t* od = A.ord + (size_t)e * A.ordStride;
        const uint16_t o0 = od[0], os = od[smallest];
        A.lohi[i0] = xs; A.lohi[is] = x0;
        od[0] = os; od[smallest] = o0;
        const bool c0 = maskGet<W>(cand, 0), cs = maskGet<W>(cand, smallest);
        maskPut<W>(cand, 0, cs); maskPut<W>(cand, smallest, c0);
        const bool v0 = maskGet<W>(vmask, 0), vs = maskGet<W>(vmask, smallest);
        maskPut<W>(vmask, 0, vs); maskPut<W>(vmask, smallest, v0);
      }
      // progressive.cc:104-117
      const double tprev = t;
      t = reverse ? ilo : ihi;
      lastValid = tprev;
      hp = A.alpha * fabs(t - tprev);
      if (reverse ? t <= 0.0 : t >= L) { t = reverse ? 0.0 : L; finished++; }
      if (finished >= 2) finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
      else {
        needSample = true;
        // the first non-valid pair of the new order is evaluated by the reference for sure
        int first = -1;
#pragma unroll
        for (int w = W - 1; w >= 0; --w) {
          const u64 nv = ~vmask[w];
          if (nv) first = w * 64 + __ffsll((long long)nv) - 1;
        }
        if (first >= 0 && first < NP) maskPut<W>(cand, first, true);
#pragma unroll
        for (int w = 0; w < W; ++w) cand[w] &= ~vmask[w];
      }
    }
  }

  // ---- REFILL: free slots pull the next edges (ballot + popc prefix, one atomic per warp) --------  @region refill
  {
    const bool want = !active && !retired;
    const unsigned m = __ballot_sync(FULL, want);
    if (m) {
      const int leader = __ffs(m) - 1;
      unsigned long long b = 0;
      if (lane == leader) b = atomicAdd(A.counter, (unsigned long long)__popc(m));
      b = __shfl_sync(FULL, b, leader);
      bool got = false;
      if (want) {
        const unsigned long long idx = b + __popc(m & ((1u << lane) - 1u));
        if (idx >= A.n) retired = true;
        else {
          edge = A.perm ? A.perm[idx] : (uint32_t)idx;
          const double* q1 = A.q1 + (size_t)edge * mv.nq;
          const double* q2 = A.q2 + (size_t)edge * mv.nq;
          L = A.L[edge];
          iters = 0;
          finished = 0;
          lastValid = reverse ? L : 0.0;
          t = lastValid;
          hp = INF;
          if (reverse ? t <= 0.0 : t >= L) finished++;
#pragma unroll
          for (int w = 0; w < W; ++w) {
            vmask[w] = 0ull;
            const int left = NP - 64 * w;
            cand[w] = left >= 64 ? ~0ull : (left > 0 ? (1ull << left) - 1ull : 0ull);
          }
          active = true;
          needSample = true;
          got = true;
        }
      }
      wfNewEdgeTables(mv, A, T, e, lane, got, edge, L);
      __syncwarp();
    }
  }

  // ---- next sample ---------------------------------------------------------------------------------  @region sample
  if (needSample) {
    if (iters >= A.max_iter) {
      ++caps;
      finishEdge(false, HPPB200_REASON_ITER_CAP, t, lastValid, -1);
    } else {
      ++iters;
      ++samples;
      pos = 0; smallest = 0; ilo = -INF; ihi = INF; nEm = 0;
      needFK = true;  // wfFkKernel computes the placements at t
    }
  }

  // ---- EMIT: one job per geometry pair of every candidate, in position order -------------------------  @region emit
  wfEmitJobs<W>(A, T, e, lane, active, cand, nEm, sCnt, ctl);

  // ---- store the slot state -------------------------------------------------------------------------  @region store
  A.st_flags[e] = (uint8_t)((active ? 1 : 0) | (retired ? 2 : 0) | (finished << 2) | ((active && needFK) ? 16 : 0));
  if (active) {
    A.st_t[e] = t; A.st_last[e] = lastValid; A.st_hp[e] = hp; A.st_ilo[e] = ilo; A.st_ihi[e] = ihi; A.st_L[e] = L;
    A.st_edge[e] = edge; A.st_iters[e] = iters;
    A.st_pos[e] = (uint32_t)pos | ((uint32_t)smallest << 16);
    A.st_nEm[e] = (uint8_t)nEm;
#pragma unroll
    for (int w = 0; w < W; ++w) A.st_vmask[(size_t)w * E + e] = vmask[w];
  }
  wfPublish(A, stats, sCnt, lane, retired, samples, discarded, caps, overflows);
}

// ------------------------------------------------------------------------------------------------
// Control kernel, Progressive, second form (round 2).  Same replay, same results; what changed is HOW a thread
// walks the pair positions.  The first form walked them one by one with the whole decision tree inside the loop:
// ~12 k warp instructions per warp at 14.7 of 32 lanes (ncu r1j), because the ~5 positions per sample where the
// reference really evaluates a pair ("events") sit at different positions in different lanes and every one of
// them serialised the warp.  Here the walk is split into phases every lane runs in step:
//   1. FIRST EVENT: the first non-valid_ position from `pos` (it is evaluated for sure: the running interval is
//      still infinite, or the replay stopped there last round).
//   2. UNIFORM PASS over all later positions, 8 independent 16-byte loads in flight, no decision inside: two bit
//      masks, M = "not covered by the running interval as it is NOW" and C = "candidate of the next sample under
//      the window predicted NOW".  The running interval only shrinks, so a position covered now stays covered:
//      M and C are supersets of what the exact replay / the final prediction will need.
//   3. EVENT LOOP over the bits of M in position order (a handful): re-test against the current running interval,
//      fold the pair's result (BodyPairCollision::validateConfiguration :65-89, formerly wfFoldKernel) and
//      intersect, exactly like validateIntervals.  Lanes process their j-th event together.
//   4. Sample complete: the bits of C are re-tested against the final window; or a result was missing: the
//      remaining bits of M are the candidates of the current sample.
template <int W>
__device__ __forceinline__ int maskFirstFrom(const u64 (&m)[W], int from) {  // first set bit at a position >= from, -1 if none
#pragma unroll
  for (int w = 0; w < W; ++w) {
    const int lo = from - 64 * w;
    if (lo >= 64) continue;
    u64 x = m[w];
    if (lo > 0) x &= ~0ull << lo;
    if (x) return 64 * w + __ffsll((long long)x) - 1;
  }
  return -1;
}
template <int W>
__device__ __forceinline__ void maskClear(u64 (&m)[W], int p) {
  if (W == 1) { m[0] &= ~(1ull << p); return; }
#pragma unroll
  for (int w = 0; w < W; ++w)
    if ((p >> 6) == w) m[w] &= ~(1ull << (p & 63));
}
template <int W, int N>
__device__ __forceinline__ void chunkOrN(u64 (&m)[W], int c0, unsigned bits) {
  if (W == 1) { if (c0 < 64) m[0] |= (u64)bits << c0; return; }
#pragma unroll
  for (int k = 0; k < N; ++k) maskSet<W>(m, c0 + k, (bits >> k) & 1u);
}
constexpr int WP_CH = 8;  // positions per trip of the uniform pass
constexpr int WF_RES_STAGED = 8;  // control form 3: results per slot staged in shared memory (8 keeps 8 CTAs per SM on UR5-class models)

#ifndef WF_PC_MINB
#define WF_PC_MINB 8
#endif
// AOS (control form 3): the emission records and last round's distances of the slot come slot-major (WFArgs::emA, resA)
// and are pulled into shared memory by the thread itself with 16-byte cp.async while the blob is staged; the fold of a
// distance into a validated interval happens in the event that uses it (its only remaining load, vht, is issued next to
// the event's interval load).  Why: the kernel is bound by the NUMBER of dependent global-load steps a warp goes through
// (profiles/r2c_wfProgControl_ncu_full.txt), and the up-front fold loop alone was 4 steps x max(nEm) of them.
template <int W, bool AOS>
__global__ void __launch_bounds__(WF_CTHREADS, W == 1 ? WF_PC_MINB : 4)
wfProgControlKernel(ModelView mv, WFArgs A, DeviceStats* stats) {
  // shared memory: [0,8) mbarrier | [16,48) block counters | model blob prefix | AOS: staging rows, 16-byte chunk c of thread t at (c * 128 + t) * 16
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  unsigned* sCnt = (unsigned*)(smem + 16);
  unsigned char* sblob = smem + 48;
  unsigned char* stg = sblob + ((mv.edge_blob_bytes + 15u) & ~15u) + (size_t)threadIdx.x * 16;
  const int NE = AOS ? (A.rb * 4 + 15) / 16 : 0;  // chunks of the emission row; the result row follows
  if (AOS) {
    const uint32_t e0 = blockIdx.x * WF_CTHREADS + threadIdx.x;
    const int NR = WF_RES_STAGED / 2;  // the first WF_RES_STAGED results of the slot; later ones (rare) are read in place
    const unsigned char* ge = (const unsigned char*)(A.emA + (size_t)e0 * A.rb);
    const unsigned char* gr = (const unsigned char*)(A.resA + (size_t)e0 * A.rg);
    for (int c = 0; c < NE; ++c) cpAsync16(stg + (size_t)c * WF_CTHREADS * 16, ge + c * 16);
    for (int c = 0; c < NR; ++c) cpAsync16(stg + (size_t)(NE + c) * WF_CTHREADS * 16, gr + c * 16);
    cpAsyncCommit();
  }
  auto emAt = [&](int k) -> uint32_t { return ((const uint32_t*)(stg + (size_t)(k >> 2) * WF_CTHREADS * 16))[k & 3]; };
  auto resAt = [&](int r) -> double {
    if (r >= WF_RES_STAGED) return A.resA[(size_t)(blockIdx.x * WF_CTHREADS + threadIdx.x) * A.rg + r];
    return ((const double*)(stg + (size_t)(NE + (r >> 1)) * WF_CTHREADS * 16))[r & 1];
  };
  if (threadIdx.x < 8) sCnt[threadIdx.x] = 0u;
  stageBlobTMA(sblob, mv.blob, mv.edge_blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);

  const unsigned FULL = 0xffffffffu;
  const uint32_t E = A.E;
  const uint32_t e = blockIdx.x * WF_CTHREADS + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const int NP = mv.nbodypairs;
  const double INF = CUDART_INF;
  const bool reverse = A.reverse != 0;
  WFRoundCtl* ctl = A.ctl + (A.round & 1);
  if (e == 0) {  // the other parity was consumed by the previous round's narrow phase
    WFRoundCtl* nx = A.ctl + ((A.round + 1) & 1);
    nx->nG = 0; nx->nC = 0; nx->headG = 0; nx->headC = 0;
  }

  const uint8_t flags0 = A.st_flags[e];
  bool active = (flags0 & 1) != 0, retired = (flags0 & 2) != 0;
  int finished = (flags0 >> 2) & 3;
  double t = 0, lastValid = 0, hp = INF, ilo = -INF, ihi = INF, L = 0;
  uint32_t edge = 0, iters = 0;
  int pos = 0, smallest = 0, nEm = 0;
  if (active) {
    t = A.st_t[e]; lastValid = A.st_last[e]; hp = A.st_hp[e]; ilo = A.st_ilo[e]; ihi = A.st_ihi[e]; L = A.st_L[e];
    edge = A.st_edge[e]; iters = A.st_iters[e];
    const uint32_t ps = A.st_pos[e];
    pos = (int)(ps & 0xffffu); smallest = (int)(ps >> 16);
    nEm = A.st_nEm[e];
  }
  unsigned samples = 0, caps = 0, overflows = 0, discarded = 0;
  bool needSample = false, needFK = false;
  u64 cand[W], vmask[W];
#pragma unroll
  for (int w = 0; w < W; ++w) { cand[w] = 0ull; vmask[w] = active ? A.st_vmask[(size_t)w * E + e] : 0ull; }

  auto finishEdge = [&](bool ok, int rs, double tf, double tl, int pr) {
    A.valid[edge] = ok ? 1 : 0;
    if (A.t_fail) A.t_fail[edge] = tf;
    if (A.t_last) A.t_last[edge] = tl;
    if (A.pair) A.pair[edge] = pr;
    if (A.reason) A.reason[edge] = (uint8_t)rs;
    active = false;
  };

  // ---- FOLD: last round's distances of every emitted body pair -> validated interval, all of them up front (the loads
  // of the nEm entries are independent; inside the event loop they were a chain of four dependent global loads each).
  // BodyPairCollision::validateConfiguration :65-89 up to the interval [t - (hd + ht), t + (hd + ht)]; a collision is
  // kept as (geometry pair, -inf).  Whether an interval is USED is decided by the replay below.
  if (!AOS && active) {
    const uint16_t* od = A.ord + (size_t)e * A.ordStride;
    for (int k = 0; k < nEm; ++k) {
      const int p = A.emPos[(size_t)k * E + e];
      const int r0 = A.emRes[(size_t)k * E + e];
      const int id = od[p];
      const DevBodyPair& B = T.bodypairs[id];
      double dlb = INF;  // computeDistanceLowerBound (:175-198): geometry pairs in order, first collision reports
      int hit = -1;
      for (int gp = B.pair_begin; gp < B.pair_end; ++gp) {
        const double d = A.res[(size_t)(r0 + gp - B.pair_begin) * E + e];
        if (d <= HPPB200_COLLISION_THRESHOLD) { hit = gp; break; }
        if (d < dlb) dlb = d;
      }
      double2 f;
      if (hit >= 0) f = make_double2((double)hit, -INF);
      else {
        double hd, ht;  // :72-85
        if (dlb == INF) { hd = INF; ht = 0; }
        else {
          const double2 vh = A.vht[(size_t)id * E + e];
          hd = dlb / vh.x;
          ht = vh.y;
        }
        f = make_double2(t - (hd + ht), t + (hd + ht));
      }
      A.fold[(size_t)k * E + e] = f;
    }
  }

  if (AOS) cpAsyncWait<0>();  // (every thread reads only the chunks it copied itself)
  // AOS: BodyPairCollision::validateConfiguration :65-89 for the c-th emitted body pair, at the moment the replay uses it
  auto foldAt = [&](int c, double2 vh) -> double2 {
    const uint32_t em = emAt(c);
    const int r0 = (int)((em >> 8) & 0xffu), ng = (int)(em >> 24);
    double dlb = INF;  // computeDistanceLowerBound (:175-198): geometry pairs in order, first collision reports
    int hit = -1;
    for (int j = 0; j < ng; ++j) {
      const double d = resAt(r0 + j);
      if (d <= HPPB200_COLLISION_THRESHOLD) { hit = j; break; }
      if (d < dlb) dlb = d;
    }
    if (hit >= 0) return make_double2((double)(T.bodypairs[(em >> 16) & 0xffu].pair_begin + hit), -INF);
    double hd, ht;  // :72-85
    if (dlb == INF) { hd = INF; ht = 0; }
    else { hd = dlb / vh.x; ht = vh.y; }
    return make_double2(t - (hd + ht), t + (hd + ht));
  };
  auto posAt = [&](int c) -> int { return AOS ? (int)(emAt(c) & 0xffu) : (int)A.emPos[(size_t)c * E + e]; };
  // The phases below are separated by __syncwarp(): the event loops have data-dependent trip counts, and without
  // an explicit reconvergence point the lanes that leave a loop early run the rest of the kernel on their own
  // (measured: 3.2 of 32 lanes in the completion code).
  __syncwarp();
  double2* col = A.lohi + e;  // position p at col[p * E]
  u64 nonv[W], M[W];
#pragma unroll
  for (int w = 0; w < W; ++w) {
    const int left = NP - 64 * w;
    nonv[w] = ~vmask[w] & (left >= 64 ? ~0ull : (left > 0 ? (1ull << left) - 1ull : 0ull));
    M[w] = 0ull;
  }
  // first valid_ position from `pos`: the only one that can still change the running interval (positions before
  // `pos` were walked in an earlier round of this sample; some of them became valid_ IN this sample, as events)
  const int pv = maskFirstFrom<W>(vmask, pos);
  const bool scanning = active;  // slots that hold an edge at the start of the round
  bool vdone = false, stop = false, emit = false;
  int cursor = 0;
  int nextP = (scanning && nEm > 0) ? posAt(0) : -1;

  // events of the positions in Mk, in order (validateIntervals, continuous-validation.hh:267-298)
  auto runEvents = [&](u64 (&Mk)[W]) {
    for (;;) {
      const int p = maskFirstFrom<W>(Mk, 0);
      if (p < 0) break;
      maskClear<W>(Mk, p);
      if (!vdone && pv >= 0 && pv < p) {  // a valid_ pair before p: interval = timeRange (body-pair-collision.cc:51-55)
        if (ihi > L) smallest = pv;
        ilo = (ilo < 0.0) ? 0.0 : ilo;
        ihi = (L < ihi) ? L : ihi;
        vdone = true;
      }
      double2 lh, rr;
      if (AOS) {
        // results of pairs the replay skipped (the walk order is fixed: this does not depend on position p's interval)
        while (nextP >= 0 && nextP < p) {
          ++discarded;
          ++cursor;
          nextP = cursor < nEm ? posAt(cursor) : -1;
        }
        double2 vh = make_double2(1.0, 0.0);
        if (nextP == p) vh = A.vht[(size_t)((emAt(cursor) >> 16) & 0xffu) * E + e];  // in flight next to the interval
        lh = col[(size_t)p * E];
        if (lh.x <= ilo && ihi <= lh.y) continue;  // validInterval_ contains the running interval (:56-63)
        if (nextP != p) {  // the reference evaluates this pair now and its distance is not there yet
          emit = true;
          pos = p;
          stop = true;
          break;
        }
        rr = foldAt(cursor, vh);
        ++cursor;
        nextP = cursor < nEm ? posAt(cursor) : -1;
      } else {
      lh = col[(size_t)p * E];
      if (lh.x <= ilo && ihi <= lh.y) continue;  // validInterval_ contains the running interval (:56-63)
      while (nextP >= 0 && nextP < p) {  // results of pairs the replay skipped
        ++discarded;
        ++cursor;
        nextP = cursor < nEm ? (int)A.emPos[(size_t)cursor * E + e] : -1;
      }
      if (nextP != p) {  // the reference evaluates this pair now and its distance is not there yet
        emit = true;
        pos = p;
        stop = true;
        break;
      }
      rr = A.fold[(size_t)cursor * E + e];
      ++cursor;
      nextP = cursor < nEm ? (int)A.emPos[(size_t)cursor * E + e] : -1;
      }
      const double a = rr.x, bb = rr.y;
      if (bb == -INF) {  // collision: a carries the geometry pair
        finishEdge(false, HPPB200_REASON_COLLISION, t, lastValid, (int)a);
        stop = true;
        break;
      }
      const double l = lh.x, h = lh.y;
      if (!(l > h) && (h < a || bb < l)) {
        // validInterval_.insert (:90) would open a second interval: hand the edge to the walker
        ++overflows;
        finishEdge(false, A.overflow_reason, t, lastValid, -1);
        stop = true;
        break;
      }
      double nl = a, nh = bb;
      if (!(l > h)) { nl = (a < l) ? a : l; nh = (h > bb) ? h : bb; }
      if (nl <= 0.0 && L <= nh) {  // valid_ (:92-96), stored as lo = -inf
        nl = -INF; nh = INF;
        maskSet<W>(vmask, p, true);
        maskClear<W>(nonv, p);
      }
      col[(size_t)p * E] = make_double2(nl, nh);
      if (ihi > bb) smallest = p;  // validateIntervals: fold into the running intersection
      ilo = (ilo < a) ? a : ilo;
      ihi = (bb < ihi) ? bb : ihi;
    }
  };
  // uniform pass over the positions of `need`: Mo = those whose interval does not contain [xl, xh]
  auto pass = [&](const u64 (&need)[W], double xl, double xh, u64 (&Mo)[W]) {
    for (int c0 = 0; c0 < NP; c0 += WP_CH) {
      const unsigned nb = chunkBitsN<W, WP_CH>(need, c0);
      double2 v[WP_CH];
#pragma unroll
      for (int k = 0; k < WP_CH; ++k) {
        v[k] = make_double2(-INF, INF);
        if ((nb >> k) & 1u) v[k] = col[(size_t)(c0 + k) * E];
      }
      unsigned mb = 0u;
#pragma unroll
      for (int k = 0; k < WP_CH; ++k) mb |= (!(v[k].x <= xl && xh <= v[k].y) ? 1u : 0u) << k;
      chunkOrN<W, WP_CH>(Mo, c0, mb & nb);
    }
  };
  // the same with a second test in the same loads: Co = the positions whose interval does not contain [cl, ch]
  auto pass2 = [&](const u64 (&need)[W], double xl, double xh, u64 (&Mo)[W], double cl, double ch, u64 (&Co)[W]) {
    for (int c0 = 0; c0 < NP; c0 += WP_CH) {
      const unsigned nb = chunkBitsN<W, WP_CH>(need, c0);
      double2 v[WP_CH];
#pragma unroll
      for (int k = 0; k < WP_CH; ++k) {
        v[k] = make_double2(-INF, INF);
        if ((nb >> k) & 1u) v[k] = col[(size_t)(c0 + k) * E];
      }
      unsigned mb = 0u, cb = 0u;
#pragma unroll
      for (int k = 0; k < WP_CH; ++k) {
        mb |= (!(v[k].x <= xl && xh <= v[k].y) ? 1u : 0u) << k;
        cb |= (!(v[k].x <= cl && ch <= v[k].y) ? 1u : 0u) << k;
      }
      chunkOrN<W, WP_CH>(Mo, c0, mb & nb);
      chunkOrN<W, WP_CH>(Co, c0, cb & nb);
    }
  };
  auto fromMask = [&](int from, u64 (&need)[W]) {  // the non-valid_ positions >= from
#pragma unroll
    for (int w = 0; w < W; ++w) {
      const int lo = from - 64 * w;
      need[w] = lo >= 64 ? 0ull : (lo > 0 ? nonv[w] & (~0ull << lo) : nonv[w]);
    }
  };

  // ---- phase 1: the first event ----------------------------------------------------------------------
  int from = NP;
  if (scanning) {
    const int p0 = maskFirstFrom<W>(nonv, pos);
    if (p0 >= 0) {
      from = p0 + 1;
      maskSet<W>(M, p0, true);
      runEvents(M);
    }
  }
  __syncwarp();
  // ---- phase 2: uniform pass against the running interval as it is now ------------------------------------
  // The same loads also answer most of phase 4 (the candidates of the NEXT sample, window [tn - hn, tn + hn] with
  // tn = the end of the FINAL intersection and hn = alpha |tn - t|): tn can only move from the current end towards t,
  // so every such window lies inside the hull [cl, ch] computed here, and a position whose interval contains the hull
  // contains the window.  Phase 4 then re-reads only the positions flagged here (C1), the positions the events rewrote
  // and those before `from` (55 % of the kernel's sectors were the two full passes).
  const bool passed = scanning && !stop;
  u64 C1[W];
#pragma unroll
  for (int w = 0; w < W; ++w) C1[w] = 0ull;
  double hullLo = INF, hullHi = -INF;  // (empty: no narrowed second pass)
  if (passed) {
    const double te = reverse ? ilo : ihi;            // the end the parameter moves to, as it is now
    const double he = A.alpha * fabs(te - t);
    // windows of every tn between t and te (piecewise linear in tn), and the running interval itself: a position the
    // events rewrite (it does not contain the running interval) is then flagged in C1 as well
    hullLo = fmin(fmin(te - he, t), ilo) - 1e-9;
    hullHi = fmax(fmax(te + he, t), ihi) + 1e-9;
    u64 need[W];
    fromMask(from, need);
    pass2(need, ilo, ihi, M, hullLo, hullHi, C1);
  }
  __syncwarp();
  // ---- phase 3: the other events -------------------------------------------------------------------------
  if (passed) runEvents(M);
  __syncwarp();
  // ---- phase 4: candidates.  A result was missing: the later pairs the window of the CURRENT sample does not cover
  // (and the missing one); the sample is complete: the pairs the window of the NEXT sample does not cover
  // (predicted t' = end of the intersection, window alpha x step).
  const bool complete = scanning && !stop && !emit;
  double wl = 0, wh = 0;
  if (emit) { wl = fmax(ilo, t - hp); wh = fmin(ihi, t + hp); }
  if (complete) {
    if (!vdone && pv >= 0) {  // a valid_ pair after the last event
      if (ihi > L) smallest = pv;
      ilo = (ilo < 0.0) ? 0.0 : ilo;
      ihi = (L < ihi) ? L : ihi;
    }
    discarded += (unsigned)(nEm - cursor);
    const double tn = reverse ? ilo : ihi;
    const double hn = A.alpha * fabs(tn - t);
    wl = tn - hn; wh = tn + hn;
  }
  if (complete || (emit && !passed)) {
    u64 need[W];
    fromMask(complete ? 0 : pos + 1, need);
    if (complete && hullLo <= wl && wh <= hullHi) {  // narrowed: flagged by the first pass (C1), or before `from`
#pragma unroll
      for (int w = 0; w < W; ++w) {
        const int lo = from - 64 * w;
        const u64 before = lo >= 64 ? ~0ull : (lo > 0 ? ((1ull << lo) - 1ull) : 0ull);
        need[w] &= C1[w] | before;
      }
    }
    pass(need, wl, wh, cand);
  }
  __syncwarp();
  if (emit) {
    if (passed) {
      for (int left = A.rb - 1; left > 0;) {  // the remaining bits of M are after `pos`; at most rb body pairs are emitted
        const int p = maskFirstFrom<W>(M, 0);
        if (p < 0) break;
        maskClear<W>(M, p);
        const double2 lh = col[(size_t)p * E];
        if (!(lh.x <= wl && wh <= lh.y)) { maskSet<W>(cand, p, true); --left; }
      }
    }
    maskSet<W>(cand, pos, true);
  }
  __syncwarp();
  // ---- phase 5: end of the sample ---------------------------------------------------------------------
  if (complete) {
    // src/continuous-validation.cc:144-146: smallest interval first
    if (NP > 1 && smallest != 0) {
      const size_t is = (size_t)smallest * E;
      const double2 x0 = col[0], xs = col[is];
      uint16_t* odw = A.ord + (size_t)e * A.ordStride;
      const uint16_t o0 = odw[0], os = odw[smallest];
      col[0] = xs; col[is] = x0;
      odw[0] = os; odw[smallest] = o0;
      const bool c0 = maskGet<W>(cand, 0), cs = maskGet<W>(cand, smallest);
      maskPut<W>(cand, 0, cs); maskPut<W>(cand, smallest, c0);
      const bool v0 = maskGet<W>(vmask, 0), vs = maskGet<W>(vmask, smallest);
      maskPut<W>(vmask, 0, vs); maskPut<W>(vmask, smallest, v0);
    }
    // progressive.cc:104-117
    const double tprev = t;
    t = reverse ? ilo : ihi;
    lastValid = tprev;
    hp = A.alpha * fabs(t - tprev);
    if (reverse ? t <= 0.0 : t >= L) { t = reverse ? 0.0 : L; finished++; }
    if (finished >= 2) finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
    else {
      needSample = true;
      // the first non-valid pair of the new order is evaluated by the reference for sure
      int first = -1;
#pragma unroll
      for (int w = W - 1; w >= 0; --w) {
        const u64 nv = ~vmask[w];
        if (nv) first = w * 64 + __ffsll((long long)nv) - 1;
      }
      if (first >= 0 && first < NP) maskPut<W>(cand, first, true);
#pragma unroll
      for (int w = 0; w < W; ++w) cand[w] &= ~vmask[w];
    }
  }
  __syncwarp();

  // ---- REFILL: free slots pull the next edges (ballot + popc prefix, one atomic per warp) --------
  {
    const bool want = !active && !retired;
    const unsigned m = __ballot_sync(FULL, want);
    if (m) {
      const int leader = __ffs(m) - 1;
      unsigned long long b = 0;
      if (lane == leader) b = atomicAdd(A.counter, (unsigned long long)__popc(m));
      b = __shfl_sync(FULL, b, leader);
      bool got = false;
      if (want) {
        const unsigned long long idx = b + __popc(m & ((1u << lane) - 1u));
        if (idx >= A.n) retired = true;
        else {
          edge = A.perm ? A.perm[idx] : (uint32_t)idx;
          L = A.L[edge];
          iters = 0;
          finished = 0;
          lastValid = reverse ? L : 0.0;
          t = lastValid;
          hp = INF;
          if (reverse ? t <= 0.0 : t >= L) finished++;
#pragma unroll
          for (int w = 0; w < W; ++w) {
            vmask[w] = 0ull;
            const int left = NP - 64 * w;
            cand[w] = left >= 64 ? ~0ull : (left > 0 ? (1ull << left) - 1ull : 0ull);
          }
          active = true;
          needSample = true;
          got = true;
        }
      }
      wfNewEdgeTables(mv, A, T, e, lane, got, edge, L);
      __syncwarp();
    }
  }

  // ---- next sample ---------------------------------------------------------------------------------
  bool handed = false;
  if (needSample) {
    if (iters >= A.max_iter) {
      ++caps;
      finishEdge(false, HPPB200_REASON_ITER_CAP, t, lastValid, -1);
    } else if (A.handover) {
      // thin tail of the batch: the edge continues in the CTA-per-edge kernel (kernels_tail.cu) from the state stored
      // below; a round of three launches for a handful of slots is latency, not throughput
      handed = true;
      active = false;
      retired = true;
    } else {
      ++iters;
      ++samples;
      pos = 0; smallest = 0; ilo = -INF; ihi = INF; nEm = 0;
      needFK = true;  // wfFkKernel computes the placements at t
    }
  }

  // ---- EMIT: one job per geometry pair of every candidate, in position order -------------------------
  wfEmitJobs<W>(A, T, e, lane, active, cand, nEm, sCnt, ctl);

  // ---- store the slot state -------------------------------------------------------------------------
  A.st_flags[e] = (uint8_t)((active ? 1 : 0) | (retired ? 2 : 0) | (finished << 2) | ((active && needFK) ? 16 : 0));
  if (active || handed) {
    A.st_t[e] = t; A.st_last[e] = lastValid; A.st_hp[e] = hp; A.st_ilo[e] = ilo; A.st_ihi[e] = ihi; A.st_L[e] = L;
    A.st_edge[e] = edge; A.st_iters[e] = iters;
    A.st_pos[e] = (uint32_t)pos | ((uint32_t)smallest << 16);
    A.st_nEm[e] = (uint8_t)nEm;
#pragma unroll
    for (int w = 0; w < W; ++w) A.st_vmask[(size_t)w * E + e] = vmask[w];
  }
  if (handed) A.hoList[atomicAdd(A.hoCount, 1u)] = e;
  wfPublish(A, stats, sCnt, lane, retired, samples, discarded, caps, overflows);
}

// ------------------------------------------------------------------------------------------------
// Control kernel for continuousValidation::Dichotomy (reference src/continuous-validation/dichotomy.cc:68-132).
// Same round structure as Progressive; what differs:
//  * t jumps (middle of the first non-validated gap), so a body pair's validated set is a LIST of disjoint closed
//    intervals (boost::icl::interval_set in the reference): up to WD_K per pair, position-major like `lohi`.
//    Bisection towards the validated start leaves about log2(L / interval width) disjoint intervals on the
//    binding pairs before they merge (6-8 on random UR5 edges, although 99 % of the FINAL lists have <= 4); an
//    insert that needs more than WD_K hands the edge to the walker, which has 255-entry lists.  List lengths
//    are 4 bit planes kept in registers, so the scan knows which intervals to load without loading a count.
//  * the edge's valid subset (Intervals::unionInterval, src/continuous-validation/intervals.hh) is a list of up to
//    WD_KS intervals per slot.
//  * the next parameter is known only after the scan, so the candidates of the next sample come from a second,
//    emit-only scan in the same round (window = alpha x the half-width just validated).
// out-of-line copies for the Dichotomy control kernel (119 KB of SASS with everything inline: 6.4 'no instruction' stalls per issue)
static __device__ __noinline__ bool wdUnionNI(double2* v, size_t stride, int* cnt, int cap, double a, double b) {
  int c = *cnt;
  const bool ok = wdUnion(v, stride, c, cap, a, b);
  *cnt = c;
  return ok;
}
static __device__ __noinline__ bool wdContainsNI(const double2* v, size_t stride, int cnt, double a, double b) { return wdContains(v, stride, cnt, a, b); }
constexpr int WD_K = 12;
constexpr int WD_KS = 16;
constexpr int WD_CH = 8;  // first intervals of this many positions are loaded ahead of their use

#ifndef WF_DC_MINB
#define WF_DC_MINB 5
#endif
template <int W>
__global__ void __launch_bounds__(WF_CTHREADS, W == 1 ? WF_DC_MINB : 1)
wfDichControlKernel(ModelView mv, WFArgs A, DeviceStats* stats) {
  extern __shared__ __align__(16) unsigned char smem[];  // @region setup
  uint64_t* bar = (uint64_t*)smem;
  unsigned* sCnt = (unsigned*)(smem + 16);
  unsigned char* sblob = smem + 48;
  if (threadIdx.x < 8) sCnt[threadIdx.x] = 0u;
  stageBlobTMA(sblob, mv.blob, mv.edge_blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);

  const unsigned FULL = 0xffffffffu;
  const uint32_t E = A.E;
  const uint32_t e = blockIdx.x * WF_CTHREADS + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const int NP = mv.nbodypairs;
  const double INF = CUDART_INF;
  const bool reverse = A.reverse != 0;
  WFRoundCtl* ctl = A.ctl + (A.round & 1);
  if (e == 0) {
    WFRoundCtl* nx = A.ctl + ((A.round + 1) & 1);
    nx->nG = 0; nx->nC = 0; nx->headG = 0; nx->headC = 0;
  }

  const uint8_t flags0 = A.st_flags[e];
  bool active = (flags0 & 1) != 0, retired = (flags0 & 2) != 0;
  double t = 0, hp = INF, ilo = -INF, ihi = INF, L = 0;
  uint32_t edge = 0, iters = 0;
  int pos = 0, smallest = 0, nEm = 0, subCnt = 0;
  u64 cand[W], vmask[W], c0[W], c1[W], c2[W], c3[W];  // candidates, valid_, list length bit planes (by order position)
#pragma unroll
  for (int w = 0; w < W; ++w) { cand[w] = 0ull; vmask[w] = 0ull; c0[w] = 0ull; c1[w] = 0ull; c2[w] = 0ull; c3[w] = 0ull; }
  if (active) {
    t = A.st_t[e]; hp = A.st_hp[e]; ilo = A.st_ilo[e]; ihi = A.st_ihi[e]; L = A.st_L[e];
    edge = A.st_edge[e]; iters = A.st_iters[e];
    const uint32_t ps = A.st_pos[e];
    pos = (int)(ps & 0xffffu); smallest = (int)(ps >> 16);
    nEm = A.st_nEm[e];
    subCnt = A.st_subCnt[e];
#pragma unroll
    for (int w = 0; w < W; ++w) {
      vmask[w] = A.st_vmask[(size_t)w * E + e];
      c0[w] = A.st_cbits[(size_t)(0 * W + w) * E + e];
      c1[w] = A.st_cbits[(size_t)(1 * W + w) * E + e];
      c2[w] = A.st_cbits[(size_t)(2 * W + w) * E + e];
      c3[w] = A.st_cbits[(size_t)(3 * W + w) * E + e];
    }
  }
  unsigned samples = 0, caps = 0, overflows = 0, discarded = 0;
  bool needSample = false, needFK = false;
  double2* sub = A.sub + e;  // valid subset, element j at sub[j * E]

  auto getCnt = [&](int p) -> int {
    return (maskGet<W>(c0, p) ? 1 : 0) + (maskGet<W>(c1, p) ? 2 : 0) + (maskGet<W>(c2, p) ? 4 : 0) + (maskGet<W>(c3, p) ? 8 : 0);
  };
  auto setCnt = [&](int p, int v) {
    maskPut<W>(c0, p, (v & 1) != 0); maskPut<W>(c1, p, (v & 2) != 0); maskPut<W>(c2, p, (v & 4) != 0); maskPut<W>(c3, p, (v & 8) != 0);
  };
  auto listOf = [&](int p) -> double2* { return A.ivl + (size_t)p * WD_K * E + e; };  // element j at [j * E]
  // dichotomy.cc:113-121: the valid part ends where the first (last, if reverse) validated interval ends
  auto dichLast = [&]() -> double { return reverse ? sub[(size_t)(subCnt - 1) * E].x : sub[0].y; };
  auto finishEdge = [&](bool ok, int rs, double tf, double tl, int pr) {
    A.valid[edge] = ok ? 1 : 0;
    if (A.t_fail) A.t_fail[edge] = tf;
    if (A.t_last) A.t_last[edge] = tl;
    if (A.pair) A.pair[edge] = pr;
    if (A.reason) A.reason[edge] = (uint8_t)rs;
    active = false;
  };

  // One scan over the pair positions from `start`.  emitOnly = false: REPLAY of validateIntervals with last
  // round's results, switching to EMIT at the first pair the reference evaluates whose result is missing.
  // emitOnly = true: candidates of a sample that has no results yet.  Returns true when the sample completed.
  auto scan = [&](int start, bool emitOnly) -> bool {  // @region scan
    int cursor = 0;
    int nextP = (!emitOnly && nEm > 0) ? (int)A.emPos[e] : -1;
    bool emit = emitOnly, force = emitOnly, stop = false;
    double ql = t - hp, qh = t + hp;
    for (int b0 = start; b0 < NP && !stop; b0 += WD_CH) {
      double2 first[WD_CH];
      const unsigned vch = chunkBitsN<W, WD_CH>(vmask, b0);
      const unsigned k0 = chunkBitsN<W, WD_CH>(c0, b0), k1 = chunkBitsN<W, WD_CH>(c1, b0);
      const unsigned k2 = chunkBitsN<W, WD_CH>(c2, b0), k3 = chunkBitsN<W, WD_CH>(c3, b0);
      const unsigned anyc = (k0 | k1 | k2 | k3) & ~vch;  // non-valid positions with a non-empty list
#pragma unroll
      for (int k = 0; k < WD_CH; ++k) {
        first[k] = make_double2(INF, -INF);
        if (b0 + k < NP && ((anyc >> k) & 1u)) first[k] = listOf(b0 + k)[0];
      }
#pragma unroll
      for (int k = 0; k < WD_CH; ++k) {
        const int p = b0 + k;
        if (p >= NP || stop) break;
        if ((vch >> k) & 1u) {  // valid_ on the whole path: interval = timeRange (body-pair-collision.cc:51-55)
          if (!emit) {
            if (ihi > L) smallest = p;
            ilo = (ilo < 0.0) ? 0.0 : ilo;
            ihi = (L < ihi) ? L : ihi;
          }
          continue;
        }
        int cnt = (int)(((k0 >> k) & 1u) + 2u * ((k1 >> k) & 1u) + 4u * ((k2 >> k) & 1u) + 8u * ((k3 >> k) & 1u));
        double2* lst = listOf(p);
        const double x0 = emit ? ql : ilo, x1 = emit ? qh : ihi;
        bool cont = first[k].x <= x0 && x1 <= first[k].y;  // icl::contains(validInterval_, [x0, x1]) (:56-63)
        for (int j = 1; j < cnt && !cont; ++j) {  // (loading entries 1..4 together was measured slower: 654 -> 777 us)
          const double2 v = lst[(size_t)j * E];
          cont = v.x <= x0 && x1 <= v.y;
        }
        if (emit) {
          maskSet<W>(cand, p, force || !cont);
          force = false;
          continue;
        }
        const bool have = nextP == p;
        const int kk = cursor;
        if (have) {
          ++cursor;
          nextP = cursor < nEm ? (int)A.emPos[(size_t)cursor * E + e] : -1;
        }
        if (cont) {
          if (have) ++discarded;
        } else if (have) {
          const double2 rr = A.fold[(size_t)kk * E + e];
          const double a = rr.x, bb = rr.y;
          if (bb == -INF) {  // collision: a carries the geometry pair
            finishEdge(false, HPPB200_REASON_COLLISION, t, dichLast(), (int)a);
            stop = true;
            continue;
          }
          if (!wdUnionNI(lst, E, &cnt, WD_K, a, bb)) {  // validInterval_.insert (:90) needs a longer list: walker
            ++overflows;
            finishEdge(false, A.overflow_reason, t, dichLast(), -1);
            stop = true;
            continue;
          }
          setCnt(p, cnt);
          if (wdContainsNI(lst, E, cnt, 0.0, L)) maskSet<W>(vmask, p, true);  // valid_ (:92-96)
          if (ihi > bb) smallest = p;  // validateIntervals: fold into the running intersection
          ilo = (ilo < a) ? a : ilo;
          ihi = (bb < ihi) ? bb : ihi;
        } else {
          // the reference evaluates this pair now and its distance is not there yet
          emit = true;
          pos = p;
          ql = fmax(ilo, t - hp);
          qh = fmin(ihi, t + hp);
#pragma unroll
          for (int w = 0; w < W; ++w) cand[w] = 0ull;
          maskSet<W>(cand, p, true);
        }
      }
    }
    return !emit && !stop;
  };

  // (explicit reconvergence points: the scans have data-dependent trip counts, and lanes that leave one early otherwise
  //  run the rest of the kernel on their own, see wfProgControlKernel)
  __syncwarp();
  if (active) {
    const bool complete = scan(pos, false);
    if (active && complete) {
      // src/continuous-validation.cc:144-146: smallest interval first
      if (NP > 1 && smallest != 0) {
        const int n0 = getCnt(0), ns = getCnt(smallest);
        double2* l0 = listOf(0);
        double2* ls = listOf(smallest);
        const int nm = n0 > ns ? n0 : ns;
        for (int j = 0; j < nm; ++j) {
          const double2 x0 = l0[(size_t)j * E], xs = ls[(size_t)j * E];
          l0[(size_t)j * E] = xs; ls[(size_t)j * E] = x0;
        }
        setCnt(0, ns); setCnt(smallest, n0);
        uint16_t* od = A.ord + (size_t)e * A.ordStride;
        const uint16_t o0 = od[0], os = od[smallest];
        od[0] = os; od[smallest] = o0;
        const bool v0 = maskGet<W>(vmask, 0), vs = maskGet<W>(vmask, smallest);
        maskPut<W>(vmask, 0, vs); maskPut<W>(vmask, smallest, v0);
      }
      // dichotomy.cc:96-109
      hp = A.alpha * (0.5 * (ihi - ilo));
      if (!wdUnionNI(sub, E, &subCnt, WD_KS, ilo, ihi)) {
        ++overflows;
        finishEdge(false, A.overflow_reason, t, dichLast(), -1);
      } else if (wdContainsNI(sub, E, subCnt, 0.0, L)) {
        finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
      } else {
        double t0, t1;  // middle of the first non-validated gap
        if (!reverse) { t0 = sub[0].y; t1 = subCnt > 1 ? sub[E].x : L; }
        else { t0 = sub[(size_t)(subCnt - 1) * E].x; t1 = subCnt > 1 ? sub[(size_t)(subCnt - 2) * E].y : 0.0; }
        t = .5 * (t0 + t1);
        needSample = true;
      }
    }
  }

  __syncwarp();
  // ---- REFILL ---------------------------------------------------------------------------------------  @region refill
  {
    const bool want = !active && !retired;
    const unsigned m = __ballot_sync(FULL, want);
    if (m) {
      const int leader = __ffs(m) - 1;
      unsigned long long b = 0;
      if (lane == leader) b = atomicAdd(A.counter, (unsigned long long)__popc(m));
      b = __shfl_sync(FULL, b, leader);
      bool got = false;
      if (want) {
        const unsigned long long idx = b + __popc(m & ((1u << lane) - 1u));
        if (idx >= A.n) retired = true;
        else {
          edge = A.perm ? A.perm[idx] : (uint32_t)idx;
          L = A.L[edge];
          iters = 0;
          t = reverse ? L : 0.0;
          hp = INF;
          sub[0] = make_double2(t, t);  // validSubset.unionInterval((t, t)) (dichotomy.cc:84)
          subCnt = 1;
#pragma unroll
          for (int w = 0; w < W; ++w) { vmask[w] = 0ull; c0[w] = 0ull; c1[w] = 0ull; c2[w] = 0ull; c3[w] = 0ull; }
          active = true;
          needSample = true;
          got = true;
        }
      }
      wfNewEdgeTables(mv, A, T, e, lane, got, edge, L);
    }
  }

  __syncwarp();
  // ---- next sample: iteration cap, then the candidates of a sample without results -----------------  @region sample
  if (needSample) {
    if (iters >= A.max_iter) {
      ++caps;
      finishEdge(false, HPPB200_REASON_ITER_CAP, t, dichLast(), -1);
    } else {
      ++iters;
      ++samples;
      pos = 0; smallest = 0; ilo = -INF; ihi = INF; nEm = 0;
      needFK = true;  // wfFkKernel computes the placements at t
#pragma unroll
      for (int w = 0; w < W; ++w) cand[w] = 0ull;
      scan(0, true);
    }
  }

  __syncwarp();
  // ---- EMIT ------------------------------------------------------------------------------------------  @region emit
  wfEmitJobs<W>(A, T, e, lane, active, cand, nEm, sCnt, ctl);

  // ---- store the slot state ---------------------------------------------------------------------------  @region store
  A.st_flags[e] = (uint8_t)((active ? 1 : 0) | (retired ? 2 : 0) | ((active && needFK) ? 16 : 0));
  if (active) {
    A.st_t[e] = t; A.st_hp[e] = hp; A.st_ilo[e] = ilo; A.st_ihi[e] = ihi; A.st_L[e] = L;
    A.st_edge[e] = edge; A.st_iters[e] = iters;
    A.st_pos[e] = (uint32_t)pos | ((uint32_t)smallest << 16);
    A.st_nEm[e] = (uint8_t)nEm;
    A.st_subCnt[e] = (uint8_t)subCnt;
#pragma unroll
    for (int w = 0; w < W; ++w) {
      A.st_vmask[(size_t)w * E + e] = vmask[w];
      A.st_cbits[(size_t)(0 * W + w) * E + e] = c0[w];
      A.st_cbits[(size_t)(1 * W + w) * E + e] = c1[w];
      A.st_cbits[(size_t)(2 * W + w) * E + e] = c2[w];
      A.st_cbits[(size_t)(3 * W + w) * E + e] = c3[w];
    }
  }
  wfPublish(A, stats, sCnt, lane, retired, samples, discarded, caps, overflows);
}

// ------------------------------------------------------------------------------------------------
// Fold kernel: one thread per (slot, emitted body pair).  BodyPairCollision::validateConfiguration :65-89 on
// last round's distances, up to the interval [t - (hd + ht), t + (hd + ht)]; a collision is kept as
// (geometry pair, -inf).  Whether the interval is USED is decided by the control kernel's scan.
__global__ void __launch_bounds__(256)
wfFoldKernel(ModelView mv, WFArgs A) {
  const Tables T = makeTables(mv.blob, mv);  // global copy (read-only, L1 resident)
  const uint32_t E = A.E;
  const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (size_t)A.Ecur * A.rb) return;
  const uint32_t e = (uint32_t)(idx % A.Ecur);
  const int k = (int)(idx / A.Ecur);
  if (!(A.st_flags[e] & 1) || k >= (int)A.st_nEm[e]) return;
  const double INF = CUDART_INF;
  const int p = A.emPos[(size_t)k * E + e];
  const int id = A.ord[(size_t)e * A.ordStride + p];
  const int r0 = A.emRes[(size_t)k * E + e];
  const DevBodyPair B = T.bodypairs[id];
  double dlb = INF;  // computeDistanceLowerBound (:175-198): pairs in order, first collision reports
  int hit = -1;
  for (int gp = B.pair_begin; gp < B.pair_end; ++gp) {
    const double d = A.res[(size_t)(r0 + gp - B.pair_begin) * E + e];
    if (d <= HPPB200_COLLISION_THRESHOLD) { hit = gp; break; }
    if (d < dlb) dlb = d;
  }
  double ra, rb;
  if (hit >= 0) { ra = (double)hit; rb = -INF; }
  else {
    const double t = A.st_t[e];
    double hd, ht;  // :72-85
    if (dlb == INF) { hd = INF; ht = 0; }
    else {
      const double Vm = A.vmax[(size_t)id * E + e];
      hd = dlb / Vm;
      ht = (Vm != 0) ? 2 * A.tol / Vm : INF;
    }
    ra = t - (hd + ht);
    rb = t + (hd + ht);
  }
  A.fold[(size_t)k * E + e] = make_double2(ra, rb);
}

// ------------------------------------------------------------------------------------------------
// FK kernel: one thread per slot that starts a sample this round: Path::eval + forward kinematics +
// geometry placements (ContinuousValidation::validateConfiguration, src/continuous-validation.cc:134-137).
// shared memory: [0,8) mbarrier | [16, ...) model blob prefix
// Placements leave through a per-warp shared-memory tile: every lane drops its 96-byte record, then the warp
// writes the 32 records as 16-byte pieces in record order, i.e. whole 32-byte sectors (a lane-per-record
// store would touch 32 sectors with 8 useful bytes each).
// shared memory: [0,8) mbarrier | [16, 16 + warps * 3072) tiles | model blob prefix
constexpr uint32_t WF_FK_TILE = 32u * 96u;

template <bool TREE>
__global__ void __launch_bounds__(WF_CTHREADS)
wfFkKernel(ModelView mv, WFArgs A) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  unsigned char* sblob = smem + 16 + (WF_CTHREADS / 32) * WF_FK_TILE;
  stageBlobTMA(sblob, mv.blob, mv.edge_blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t e = blockIdx.x * WF_CTHREADS + threadIdx.x;
  const bool on = (A.st_flags[e] & 16) != 0;
  const unsigned onMask = __ballot_sync(FULL, on);
  if (!onMask) return;
  const uint32_t edge = on ? A.st_edge[e] : 0u;
  const double t = on ? A.st_t[e] : 0.0, L = on ? A.st_L[e] : 0.0;
  const double* q1 = A.q1 + (size_t)edge * mv.nq;
  const double* q2 = A.q2 + (size_t)edge * mv.nq;
  // StraightPath::impl_compute: stored end points are returned bit-exactly
  int qmode;
  double u = 0;
  if (t == 0.0 || L == 0.0) qmode = 0;
  else if (t == L) qmode = 1;
  else { qmode = 2; u = (t - 0.0) / L; }
  const EdgeQ qAt{q1, q2, qmode, u};
  if (TREE) {  // tree-shaped robots read parent placements back: lane-per-record stores
    if (on) forwardKinematics<true>(T, mv, qAt, A.plc + (size_t)e * A.fkSlots * 12, 1);
    return;
  }
  double2* tile = (double2*)(smem + 16 + warp * WF_FK_TILE);
  double2* gdst = (double2*)(A.plc + (size_t)(e - lane) * A.fkSlots * 12);  // the warp's 32 records are contiguous
  const size_t recStride = (size_t)A.fkSlots * 6;                           // in double2
  auto put = [&](int g, const T3& X) {  // all lanes call this in step
    double2* mine = tile + lane * 6;
    mine[0] = make_double2(X.r[0], X.r[1]); mine[1] = make_double2(X.r[2], X.r[3]); mine[2] = make_double2(X.r[4], X.r[5]);
    mine[3] = make_double2(X.r[6], X.r[7]); mine[4] = make_double2(X.r[8], X.t.x); mine[5] = make_double2(X.t.y, X.t.z);
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const int j = r * 32 + lane, rec = j / 6, part = j - rec * 6;
      if ((onMask >> rec) & 1u) gdst[(size_t)rec * recStride + (size_t)g * 6 + part] = tile[j];
    }
    __syncwarp();
  };
  // same operation sequence as forwardKinematics<false> (dev_model.cuh): a chain, joint j's parent is j - 1
  for (int g = 0; g < mv.ngeoms; ++g)
    if (T.geoms[g].joint == 0) put(g, load12(T.geoms[g].P));
  T3 cur = identityT();
  for (int j = 1; j <= mv.njoints; ++j) {
    const DevJoint& J = T.joints[j - 1];
    const T3 li = jointLocal(J, qAt);
    if (J.parent == 0) cur = li;
    else cur = compose(cur, li);
    for (int g = 0; g < mv.ngeoms; ++g)
      if (T.geoms[g].joint == j) put(g, compose(cur, load12(T.geoms[g].P)));
  }
}

// ------------------------------------------------------------------------------------------------
// Narrow-phase kernel: persistent grid over this round's job lists.
// shared memory: [0,8) mbarrier | [16, ...) model blob prefix
#ifndef WF_NMINB
#define WF_NMINB 3
#endif
constexpr int WF_REFILL_MIN = 32;  // idle lanes in a warp before new GJK jobs are fetched

__global__ void __launch_bounds__(WF_NTHREADS, WF_NMINB)
wfNarrowKernel(ModelView mv, WFArgs A, DeviceStats* stats) {
  extern __shared__ __align__(16) unsigned char smem[];  // @region setup
  uint64_t* bar = (uint64_t*)smem;
  unsigned char* sblob = smem + 16;
  stageBlobTMA(sblob, mv.blob, mv.edge_blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  WFRoundCtl* ctl = A.ctl + (A.round & 1);
  const unsigned nG = ctl->nG, nC = ctl->nC;
  const size_t slotStride = (size_t)A.fkSlots * 12;
  const uint32_t E = A.E;
  unsigned tests = 0, iterations = 0, capHits = 0;

  // ---- GJK jobs: one iteration per trip, finished lanes refilled from the list -------------------  @region gjk
  if (nG) {
    unsigned wcur = 0, wend = 0;
    bool exhausted = false, have = false;
    GjkState s;
    // per-thread job constants in shared memory, entry k at srel[k * WF_NTHREADS]: rel (12), the two inflations, the margin
    double* srel = (double*)(sblob + ((mv.edge_blob_bytes + 15u) & ~15u)) + threadIdx.x;
    const RelStrided<WF_NTHREADS> relv{srel};
    const DevShape* sa = nullptr;
    const DevShape* sb = nullptr;
    double* out = nullptr;
    for (;;) {
      const unsigned idle = __ballot_sync(FULL, !have);
      if (!exhausted && __popc(idle) >= A.refill_min) {
        bool need = !have;
        for (;;) {
          const unsigned m = __ballot_sync(FULL, need);
          if (!m) break;
          if (wcur >= wend) {
            unsigned b = 0;
            if (lane == 0) b = atomicAdd(&ctl->headG, (unsigned)WF_CHUNK);
            b = __shfl_sync(FULL, b, 0);
            if (b >= nG) { exhausted = true; break; }
            wcur = b;
            wend = (b + WF_CHUNK < nG) ? b + WF_CHUNK : nG;
          }
          const unsigned idx = wcur + __popc(m & ((1u << lane) - 1u));
          if (need && idx < wend) {
            const unsigned long long job = A.jobsG[idx];
            const uint32_t e = (uint32_t)job;
            const int gp = (int)((job >> 32) & 0x00ffffffu), ri = (int)(job >> 56);
            const DevPair pr = T.bgpairs[gp];
            const double* store = A.plc + (size_t)e * slotStride;
            T3 TA, TB;
            if (pr.a >= 0) { sa = (const DevShape*)&T.geoms[pr.a].shape; TA = load12(store + (size_t)pr.a * 12); }
            else { const DevObstacle& O = T.obstacles[-(pr.a + 1)]; sa = (const DevShape*)&O.shape; TA = load12(O.P); }
            if (pr.b >= 0) { sb = (const DevShape*)&T.geoms[pr.b].shape; TB = load12(store + (size_t)pr.b * 12); }
            else { const DevObstacle& O = T.obstacles[-(pr.b + 1)]; sb = (const DevShape*)&O.shape; TB = load12(O.P); }
            const T3 rel = composeInv(TA, TB);
            gjkBegin(s, rel);
#pragma unroll
            for (int k = 0; k < 9; ++k) srel[k * WF_NTHREADS] = rel.r[k];
            srel[9 * WF_NTHREADS] = rel.t.x; srel[10 * WF_NTHREADS] = rel.t.y; srel[11 * WF_NTHREADS] = rel.t.z;
            srel[12 * WF_NTHREADS] = inflation(*sa); srel[13 * WF_NTHREADS] = inflation(*sb); srel[14 * WF_NTHREADS] = pr.margin;
            out = A.resA ? A.resA + (size_t)e * A.rg + ri : A.res + (size_t)ri * E + e;
            have = true;
            need = false;
            ++tests;
          }
          const unsigned c = (unsigned)__popc(m);
          wcur = (wcur + c < wend) ? wcur + c : wend;
        }
      }
      if (!__any_sync(FULL, have)) break;  // nothing running and nothing left to fetch
      if (have) {
        double core;
        if (gjkStep(s, *sa, *sb, relv, mv.vertices, core)) {
          // shapeDistance (GJK kinds) - margin, same operation order as pairLowerBound
          *out = ((core - srel[12 * WF_NTHREADS]) - srel[13 * WF_NTHREADS]) - srel[14 * WF_NTHREADS];
          iterations += (unsigned)s.it + 1u;
          if (s.it == HPPB200_GJK_MAX_ITER) ++capHits;
          have = false;
        }
      }
    }
  }

  // ---- closed-form jobs: one thread each ----------------------------------------------------------  @region closed
  {
    GjkCounters cnt = {0u, 0u, 0u};
    for (size_t i = (size_t)blockIdx.x * WF_NTHREADS + threadIdx.x; i < nC; i += (size_t)gridDim.x * WF_NTHREADS) {
      const unsigned long long job = A.jobsC[i];
      const uint32_t e = (uint32_t)job;
      const int gp = (int)((job >> 32) & 0x00ffffffu), ri = (int)(job >> 56);
      const DevPair pr = T.bgpairs[gp];
      *(A.resA ? A.resA + (size_t)e * A.rg + ri : A.res + (size_t)ri * E + e) = pairLowerBound(T, pr, A.plc + (size_t)e * slotStride, 1, mv.vertices, cnt);
      ++tests;
    }
    iterations += cnt.iterations;
  }

  const unsigned ts = __reduce_add_sync(FULL, tests), gi = __reduce_add_sync(FULL, iterations);  // @region stats
  const unsigned gc = __reduce_add_sync(FULL, capHits);
  if (lane == 0 && stats) {
    if (ts) atomicAdd(&stats->pair_tests, (unsigned long long)ts);
    if (gi) atomicAdd(&stats->gjk_iterations, (unsigned long long)gi);
    if (gc) atomicAdd(&stats->gjk_cap_hits, (unsigned long long)gc);
  }
}

// ------------------------------------------------------------------------------------------------
__global__ void wfInitKernel(WFArgs A) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < A.E) A.st_flags[e] = 0;
  if (e == 0) {
    *A.counter = 0ull;
    for (int k = 0; k < 2; ++k) { A.ctl[k].nG = 0; A.ctl[k].nC = 0; A.ctl[k].headG = 0; A.ctl[k].headC = 0; }
  }
}

// The last few edges of a batch can need hundreds of samples more than the rest; a round of four launches
// for a handful of slots is latency, not throughput.  The host stops the rounds when few slots are left and
// this kernel hands their edges to the walker through the overflow protocol (they restart from t = 0).
__global__ void wfAbandonKernel(WFArgs A, DeviceStats* stats) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= A.E || !(A.st_flags[e] & 1)) return;
  const uint32_t edge = A.st_edge[e];
  A.valid[edge] = 0;
  if (A.reason) A.reason[edge] = (uint8_t)A.overflow_reason;
  A.st_flags[e] = 2;
  atomicAdd(&stats->overflow, 1ull);
}

// ------------------------------------------------------------------------------------------------
// Slot compaction.  Once the edge counter is exhausted, finished slots stay empty; the thread-per-slot kernels
// then pay a full warp for every warp that still holds ONE edge (measured r2: with 2 % of the slots alive the
// control kernel still took 25 % of a full round, and 720 of the 896 rounds of a 1 M-edge batch are such rounds).
// Between two rounds (job lists consumed, nothing in flight) the host moves the live slots above the new extent
// into the free slots below it and shrinks the grids: a round then costs what its live edges cost.
// Which slot an edge lives in never reaches a result.
struct WFMoveEntry {
  unsigned char* base;
  uint32_t elem;        // bytes per element
  uint32_t count;       // elements per slot
  uint32_t slotMajor;   // 1: [E][count] (contiguous per slot), 0: [count][E]
};
constexpr int WF_MOVE_MAX = 32;
struct WFMoveTable {
  WFMoveEntry ent[WF_MOVE_MAX];
  int n;
};

__global__ void wfCompactListKernel(WFArgs A, uint32_t Ecur, uint32_t Enew, uint32_t* src, uint32_t* dst, unsigned* cnt) {
  const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= Ecur) return;
  const bool live = (A.st_flags[e] & 1) != 0;
  if (e >= Enew) { if (live) src[atomicAdd(&cnt[0], 1u)] = e; }
  else if (!live) dst[atomicAdd(&cnt[1], 1u)] = e;
}

// one warp per moved slot
__global__ void __launch_bounds__(256)
wfCompactMoveKernel(WFArgs A, WFMoveTable tab, const uint32_t* src, const uint32_t* dst, const unsigned* cnt) {
  const unsigned w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= cnt[0]) return;
  const uint32_t s = src[w], d = dst[w];
  const size_t E = A.E;
  for (int i = 0; i < tab.n; ++i) {
    const WFMoveEntry en = tab.ent[i];
    if (en.slotMajor) {
      const size_t bytes = (size_t)en.elem * en.count;  // multiples of 4 by construction
      const uint32_t* ps = (const uint32_t*)(en.base + s * bytes);
      uint32_t* pd = (uint32_t*)(en.base + d * bytes);
      for (size_t k = lane; k < bytes / 4; k += 32) pd[k] = ps[k];
    } else {
      for (uint32_t k = lane; k < en.count; k += 32) {
        const unsigned char* ps = en.base + ((size_t)k * E + s) * en.elem;
        unsigned char* pd = en.base + ((size_t)k * E + d) * en.elem;
        switch (en.elem) {
          case 16: *(double2*)pd = *(const double2*)ps; break;
          case 8: *(unsigned long long*)pd = *(const unsigned long long*)ps; break;
          case 4: *(uint32_t*)pd = *(const uint32_t*)ps; break;
          case 2: *(uint16_t*)pd = *(const uint16_t*)ps; break;
          default: *pd = *ps; break;
        }
      }
    }
  }
  __syncwarp();
  if (lane == 0) A.st_flags[s] = 2;  // (its flags were copied to d above)
}

namespace {
struct Carver {
  size_t off = 0;
  template <class T>
  size_t take(size_t count) {
    off = (off + 255) & ~size_t(255);
    const size_t o = off;
    off += count * sizeof(T);
    return o;
  }
};
}  // namespace

// Returns false when the model does not fit this engine (the caller then uses another path).
bool launchWavefront(Workspace* ws, int method, double tol, int reverse, const double* d_q1, const double* d_q2,
                                const double* d_L, size_t n, uint8_t* d_valid, double* d_tfail, double* d_tlast,
                                int32_t* d_pair, uint8_t* d_reason, int overflow_reason) {
  const ModelView& mv = ws->view;
  const Model* m = ws->model;
  const int NP = mv.nbodypairs;
  if (NP < 1 || NP > 65535 || mv.nbgpairs >= (1 << 24)) return false;
  for (const BodyPairHost& b : m->bodyPairs)
    if ((int)b.pairs.size() > WF_RG) return false;  // (a body pair must fit one round's job cap)
  if (NP > 256) return false;
  const int W = NP <= 64 ? 1 : 4;
  const size_t smemC = 48 + WF_RING_BYTES + mv.edge_blob_bytes, smemN = 16 + ((mv.edge_blob_bytes + 15u) & ~15u) + 15u * WF_NTHREADS * sizeof(double);
  if (smemC > 200u * 1024u) return false;
  const bool tree = modelIsTree(m);

  WFArgs A{};
  A.q1 = d_q1; A.q2 = d_q2; A.L = d_L; A.n = n; A.reverse = reverse; A.tol = tol; A.max_iter = ws->max_iterations;
  A.valid = d_valid; A.t_fail = d_tfail; A.t_last = d_tlast; A.pair = d_pair; A.reason = d_reason;
  A.overflow_reason = overflow_reason;
  A.perm = (ws->wavefront_lpt && n > 1) ? sortEdgesByLength(ws, d_L, n) : nullptr;
  A.alpha = ws->engine_alpha;
  // Progressive, second control form: the candidate window is computed from the FINAL intersection of the sample, so it
  // needs no slack: alpha 1.0 (discarded tests 9 % -> 1.3 %, same number of rounds; 217 -> 209 ms per 1 M UR5 edges)
  if (method == HPPB200_METHOD_PROGRESSIVE) {
    A.alpha = 1.0;
    if (const char* v = std::getenv("HPPB200_WF_ALPHA")) A.alpha = std::atof(v);
  }
  A.refill_min = WF_REFILL_MIN;
  if (const char* v = std::getenv("HPPB200_WF_REFILL")) A.refill_min = std::atoi(v);
  A.fkSlots = mv.ngeoms + (tree ? mv.njoints : 0);
  size_t E = ws->wavefront_slots ? ws->wavefront_slots : (size_t)ws->num_sms * (method == HPPB200_METHOD_DICHOTOMY ? 1024 : 2048);
  if (E > n) E = n;
  E = (E + WF_CTHREADS - 1) / WF_CTHREADS * WF_CTHREADS;
  A.E = (uint32_t)E;
  const int BATCH = 16;  // rounds between two host checks of the alive counter
  // tail: at most one wave of the walker's second pass, and never more than 1/64 of the batch
  const unsigned tailSlots = ws->wavefront_tail >= 0 ? (unsigned)ws->wavefront_tail
                             : (method == HPPB200_METHOD_DICHOTOMY ? (unsigned)std::min<size_t>(2048, n / 64) : 0u);
  // Progressive: measured slower with the hand-off (262 -> 407 ms per 1M edges).  Dichotomy: a few edges in a
  // million need thousands of samples and lists beyond any fixed capacity; they end in the walker's
  // large-list pass either way (its latency, ~0.4 s for the longest edge, is the floor of a 1M batch), so the
  // thin rounds before it are cut short (886 ms with the hand-off, 1008 ms without).

  Carver c;
  const size_t o_counter = c.take<unsigned long long>(1), o_ctl = c.take<WFRoundCtl>(2), o_alive = c.take<unsigned>(BATCH);
  const bool dich = method == HPPB200_METHOD_DICHOTOMY;
  const size_t o_lohi = c.take<double2>(dich ? 0 : (size_t)NP * E), o_ord = c.take<uint16_t>((size_t)((NP + 15) / 16 * 16) * E);
  int controlForm = 3;  // Progressive: 3 = form 2 with slot-major emission / result rows prefetched to shared memory and the fold inside the events;
                        // 2 = phase-convergent control kernel with the fold up front; 1 = first form + wfFoldKernel
  if (const char* v = std::getenv("HPPB200_WF_CONTROL")) controlForm = std::atoi(v);
  const bool foldKernel = dich || controlForm == 1;
  const size_t o_vmax = c.take<double>(foldKernel ? (size_t)NP * E : 0), o_sv = c.take<double>((size_t)mv.nv * E);
  const size_t o_vht = c.take<double2>(foldKernel ? 0 : (size_t)NP * E);
  const size_t o_t = c.take<double>(E), o_last = c.take<double>(E), o_hp = c.take<double>(E), o_ilo = c.take<double>(E);
  const size_t o_ihi = c.take<double>(E), o_L = c.take<double>(E);
  const size_t o_edge = c.take<uint32_t>(E), o_iters = c.take<uint32_t>(E), o_pos = c.take<uint32_t>(E);
  const size_t o_flags = c.take<uint8_t>(E), o_nEm = c.take<uint8_t>(E), o_vmask = c.take<unsigned long long>((size_t)W * E);
  const size_t o_cbits = c.take<unsigned long long>(dich ? (size_t)4 * W * E : 0), o_ivl = c.take<double2>(dich ? (size_t)NP * WD_K * E : 0);
  const size_t o_sub = c.take<double2>(dich ? (size_t)WD_KS * E : 0), o_subCnt = c.take<uint8_t>(dich ? E : 0);
  A.rb = NP > 64 ? WF_RB_BIG : WF_RB;
  A.rg = NP > 64 ? WF_RG_BIG : WF_RG;
  const size_t RB = (size_t)A.rb, RG = (size_t)A.rg;
  const bool aos = !dich && controlForm == 3;
  const size_t o_emA = c.take<uint32_t>(aos ? RB * E : 0), o_resA = c.take<double>(aos ? RG * E : 0);
  const size_t o_emPos = c.take<uint16_t>(RB * E), o_emRes = c.take<uint8_t>(RB * E);
  const size_t o_fold = c.take<double2>(RB * E);
  const size_t o_res = c.take<double>(RG * E), o_plc = c.take<double>(E * (size_t)A.fkSlots * 12);
  const size_t o_jG = c.take<unsigned long long>(E * RG), o_jC = c.take<unsigned long long>(E * RG);
  const size_t o_cmp = c.take<uint32_t>(2 * E + 4);  // compaction: source slots, destination slots, two counters
  const size_t o_ho = c.take<uint32_t>(E + 4);       // hand-over to the CTA-per-edge kernel: counter, slot list
  ws->d_state.reserve(c.off + 256);
  unsigned char* base = ws->d_state.p;
  A.counter = (unsigned long long*)(base + o_counter);
  A.ctl = (WFRoundCtl*)(base + o_ctl);
  A.alive = (unsigned*)(base + o_alive);
  A.lohi = dich ? nullptr : (double2*)(base + o_lohi); A.ord = (uint16_t*)(base + o_ord); A.ordStride = (uint32_t)((NP + 15) / 16 * 16);
  A.vmax = (double*)(base + o_vmax); A.sv = (double*)(base + o_sv);
  A.vht = foldKernel ? nullptr : (double2*)(base + o_vht);
  A.st_t = (double*)(base + o_t); A.st_last = (double*)(base + o_last); A.st_hp = (double*)(base + o_hp);
  A.st_ilo = (double*)(base + o_ilo); A.st_ihi = (double*)(base + o_ihi); A.st_L = (double*)(base + o_L);
  A.st_edge = (uint32_t*)(base + o_edge); A.st_iters = (uint32_t*)(base + o_iters); A.st_pos = (uint32_t*)(base + o_pos);
  A.st_flags = base + o_flags; A.st_nEm = base + o_nEm; A.st_vmask = (unsigned long long*)(base + o_vmask);
  A.st_cbits = (unsigned long long*)(base + o_cbits); A.ivl = (double2*)(base + o_ivl); A.sub = (double2*)(base + o_sub);
  A.st_subCnt = base + o_subCnt;
  A.emPos = (uint16_t*)(base + o_emPos); A.emRes = base + o_emRes; A.fold = (double2*)(base + o_fold);
  A.res = (double*)(base + o_res); A.plc = (double*)(base + o_plc);
  A.emA = aos ? (uint32_t*)(base + o_emA) : nullptr;
  A.resA = aos ? (double*)(base + o_resA) : nullptr;
  A.jobsG = (unsigned long long*)(base + o_jG); A.jobsC = (unsigned long long*)(base + o_jC);

  const size_t smemCtl = (foldKernel && !dich) ? smemC
                         : 48 + ((mv.edge_blob_bytes + 15u) & ~15u) + (aos ? ((RB * 4 + 15) / 16 + WF_RES_STAGED / 2) * 16 * (size_t)WF_CTHREADS : 0);
  auto kc = dich ? (W == 1 ? wfDichControlKernel<1> : wfDichControlKernel<4>)
                 : (controlForm == 1 ? (W == 1 ? wfControlKernel<1> : wfControlKernel<4>)
                    : controlForm == 2 ? (W == 1 ? wfProgControlKernel<1, false> : wfProgControlKernel<4, false>)
                                       : (W == 1 ? wfProgControlKernel<1, true> : wfProgControlKernel<4, true>));
  auto kf = tree ? wfFkKernel<true> : wfFkKernel<false>;
  const size_t smemF = 16 + (WF_CTHREADS / 32) * WF_FK_TILE + mv.edge_blob_bytes;
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemF));
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemCtl));
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(wfNarrowKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemN));
  int perSM = 0;
  HPPB200_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, wfNarrowKernel, WF_NTHREADS, smemN));
  if (perSM < 1) return false;
  const unsigned gridN = (unsigned)(ws->num_sms * perSM);
  unsigned gridC = (unsigned)(E / WF_CTHREADS), gridF = (unsigned)((E * RB + 255) / 256);
  A.Ecur = (uint32_t)E;

  // slot compaction (see wfCompactMoveKernel): every per-slot array that carries state from one round to the next
  WFMoveTable mt{};
  auto addMove = [&](void* base, size_t elem, size_t count, bool slotMajor) {
    if (!base || !count || mt.n >= WF_MOVE_MAX) return;
    mt.ent[mt.n++] = WFMoveEntry{(unsigned char*)base, (uint32_t)elem, (uint32_t)count, slotMajor ? 1u : 0u};
  };
  addMove(A.lohi, 16, NP, false); addMove(A.ord, 2, A.ordStride, true);
  addMove(foldKernel ? (void*)A.vmax : (void*)A.vht, foldKernel ? 8 : 16, NP, false);
  addMove(A.st_t, 8, 1, false); addMove(A.st_last, 8, 1, false); addMove(A.st_hp, 8, 1, false); addMove(A.st_ilo, 8, 1, false);
  addMove(A.st_ihi, 8, 1, false); addMove(A.st_L, 8, 1, false); addMove(A.st_edge, 4, 1, false); addMove(A.st_iters, 4, 1, false);
  addMove(A.st_pos, 4, 1, false); addMove(A.st_flags, 1, 1, false); addMove(A.st_nEm, 1, 1, false);
  addMove(A.st_vmask, 8, W, false);
  if (dich) {
    addMove(A.st_cbits, 8, (size_t)4 * W, false); addMove(A.ivl, 16, (size_t)NP * WD_K, false);
    addMove(A.sub, 16, WD_KS, false); addMove(A.st_subCnt, 1, 1, false);
  }
  if (aos) { addMove(A.emA, 4, RB, true); addMove(A.resA, 8, RG, true); }
  else { addMove(A.emPos, 2, RB, false); addMove(A.emRes, 1, RB, false); addMove(A.res, 8, RG, false); }
  addMove(A.plc, 8, (size_t)A.fkSlots * 12, true);
  uint32_t* d_src = (uint32_t*)(base + o_cmp);
  uint32_t* d_dst = d_src + E;
  unsigned* d_cnt = (unsigned*)(d_dst + E);
  bool compactOn = true;
  if (const char* v = std::getenv("HPPB200_WF_COMPACT")) compactOn = std::atoi(v) != 0;
  // Progressive: when this few slots are alive the edges continue one CTA each (kernels_tail.cu), from where they are
  A.hoCount = (unsigned*)(base + o_ho);
  A.hoList = (uint32_t*)(base + o_ho) + 4;
  unsigned handSlots = (!dich && !foldKernel) ? 4096u : 0u;
  if (const char* v = std::getenv("HPPB200_WF_HAND")) handSlots = (!dich && !foldKernel) ? (unsigned)std::atol(v) : 0u;
  const int tailK = std::min(64, (int)(65536 / (16 * std::max(1, NP))));
  if (tailK < 8) handSlots = 0;
  HPPB200_CUDA_CHECK(cudaMemsetAsync(A.hoCount, 0, 4 * sizeof(unsigned), ws->stream));
  int batch = BATCH;

  wfInitKernel<<<(unsigned)((E + 255) / 256), 256, 0, ws->stream>>>(A);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  int round = 0;
  unsigned h_alive = 1;
  unsigned long long h_counter = 0;
  const bool trace = std::getenv("HPPB200_WF_TRACE") != nullptr;  // per-round kernel times on stderr (debug)
  std::vector<cudaEvent_t> evs;
  struct TraceRow { unsigned alive, nG, nC; };
  std::vector<TraceRow> traceRows;
  const bool traceAll = trace && std::atoi(std::getenv("HPPB200_WF_TRACE")) > 1;
  while (h_alive) {
    HPPB200_CUDA_CHECK(cudaMemsetAsync(A.alive, 0, BATCH * sizeof(unsigned), ws->stream));
    for (int k = 0; k < batch; ++k, ++round) {
      A.round = round;
      A.aliveIdx = k;
      auto mark = [&]() {
        if (!trace) return;
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        cudaEventRecord(ev, ws->stream);
        evs.push_back(ev);
      };
      mark();
      if (round > 0 && foldKernel) wfFoldKernel<<<gridF, 256, 0, ws->stream>>>(mv, A);
      mark();
      kc<<<gridC, WF_CTHREADS, smemCtl, ws->stream>>>(mv, A, ws->d_stats);
      mark();
      kf<<<gridC, WF_CTHREADS, smemF, ws->stream>>>(mv, A);
      mark();
      wfNarrowKernel<<<gridN, WF_NTHREADS, smemN, ws->stream>>>(mv, A, ws->d_stats);
      mark();
      if (traceAll) {  // debug only: one synchronisation per round
        TraceRow row{};
        WFRoundCtl c{};
        cudaMemcpyAsync(&c, A.ctl + (round & 1), sizeof(c), cudaMemcpyDeviceToHost, ws->stream);
        cudaMemcpyAsync(&row.alive, A.alive + k, sizeof(unsigned), cudaMemcpyDeviceToHost, ws->stream);
        cudaStreamSynchronize(ws->stream);
        row.nG = c.nG; row.nC = c.nC;
        traceRows.push_back(row);
      }
    }
    HPPB200_CUDA_CHECK(cudaGetLastError());
    ws->stats.kernel_launches += (foldKernel ? 4 : 3) * batch;
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(&h_alive, A.alive + (batch - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(&h_counter, A.counter, sizeof(h_counter), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    const bool exhausted = h_counter >= (unsigned long long)n;  // every edge has been handed to a slot
    if (h_alive && h_alive <= tailSlots) {  // (slots hold edges only: the edge counter is exhausted long before)
      wfAbandonKernel<<<(unsigned)((A.Ecur + 255) / 256), 256, 0, ws->stream>>>(A, ws->d_stats);
      HPPB200_CUDA_CHECK(cudaGetLastError());
      ws->stats.kernel_launches += 1;
      break;
    }
    if (handSlots && !A.handover && exhausted && h_alive && h_alive <= handSlots) {
      A.handover = 1;  // (slots in the middle of a sample finish it first: a few more rounds)
      batch = 2;
    }
    // fewer live slots than the grids cover (only once the edge counter is exhausted): compact and shrink
    const uint32_t Enew = (uint32_t)(((size_t)h_alive + WF_CTHREADS - 1) / WF_CTHREADS * WF_CTHREADS);
    if (compactOn && !A.handover && exhausted && h_alive && (size_t)Enew * 8 <= (size_t)A.Ecur * 7) {
      HPPB200_CUDA_CHECK(cudaMemsetAsync(d_cnt, 0, 2 * sizeof(unsigned), ws->stream));
      wfCompactListKernel<<<(A.Ecur + 255) / 256, 256, 0, ws->stream>>>(A, A.Ecur, Enew, d_src, d_dst, d_cnt);
      const unsigned maxMoves = A.Ecur - Enew;  // at most that many live slots above the new extent
      wfCompactMoveKernel<<<(maxMoves + 7) / 8, 256, 0, ws->stream>>>(A, mt, d_src, d_dst, d_cnt);
      HPPB200_CUDA_CHECK(cudaGetLastError());
      ws->stats.kernel_launches += 2;
      A.Ecur = Enew;
      gridC = Enew / WF_CTHREADS;
      gridF = (unsigned)(((size_t)Enew * RB + 255) / 256);
    }
  }
  if (A.handover) {
    unsigned nho = 0;
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(&nho, A.hoCount, sizeof(unsigned), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    if (trace) std::fprintf(stderr, "[wf] %u edges continue one CTA each after round %d\n", nho, round);
    if (nho) {
      TailResume R;
      R.slots = A.hoList; R.E = A.E; R.ordStride = A.ordStride; R.lohi = A.lohi; R.ord = A.ord; R.st_vmask = A.st_vmask;
      R.st_t = A.st_t; R.st_last = A.st_last; R.st_iters = A.st_iters; R.st_edge = A.st_edge; R.st_flags = A.st_flags;
      if (!launchEdgeTail(ws, method, tol, reverse, d_q1, d_q2, d_L, nullptr, nho, d_valid, d_tfail, d_tlast, d_pair, d_reason,
                          overflow_reason, true, &R))
        throw std::runtime_error("wavefront hand-over: the CTA-per-edge kernel does not fit this model");
    }
  }
  if (trace) {
    double tt[4] = {0, 0, 0, 0};
    const bool all = std::atoi(std::getenv("HPPB200_WF_TRACE")) > 1;  // 2: every round, with the alive slots and job counts
    for (size_t r = 0; r + 4 < evs.size(); r += 5) {
      float d[4];
      for (int j = 0; j < 4; ++j) { cudaEventElapsedTime(&d[j], evs[r + j], evs[r + j + 1]); tt[j] += d[j]; }
      const size_t k = r / 5;
      if (all && k < traceRows.size())
        std::fprintf(stderr, "[wfr] %zu %u %u %u %.1f %.1f %.1f %.1f\n", k, traceRows[k].alive, traceRows[k].nG, traceRows[k].nC,
                     d[0] * 1e3, d[1] * 1e3, d[2] * 1e3, d[3] * 1e3);
      else if (k < 4 || k % 64 == 0)
        std::fprintf(stderr, "[wf] round %zu fold %.1f control %.1f fk %.1f narrow %.1f us\n", k, d[0] * 1e3, d[1] * 1e3, d[2] * 1e3, d[3] * 1e3);
    }
    std::fprintf(stderr, "[wf] E=%u gridC=%u gridN=%u rounds=%d fold %.2f control %.2f fk %.2f narrow %.2f ms\n", A.E, gridC, gridN, round,
                 tt[0], tt[1], tt[2], tt[3]);
    for (cudaEvent_t ev : evs) cudaEventDestroy(ev);
  }
  return true;
}

}  // namespace hppb200
