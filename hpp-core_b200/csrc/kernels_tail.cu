// Dichotomy and Progressive, one CTA per edge: the pass that finishes the edges the batch engines hand over (interval
// lists beyond the wavefront's capacity, the thin tail of a batch) and the path of small batches.  METHOD selects
// the outer loop (dichotomy.cc:68-132 / progressive.cc:78-129); the sample itself (validateConfiguration) is shared.
// Progressive can RESUME edges from the wavefront engine's slot state at a sample boundary (TailResume).
//
// Replaces Dichotomy::validateStraightPath (src/continuous-validation/dichotomy.cc:68-132) +
// ContinuousValidation::validateConfiguration / validateIntervals (src/continuous-validation.cc:128-148,
// include/hpp/core/continuous-validation.hh:267-298) + BodyPairCollision::validateConfiguration
// (src/continuous-validation/body-pair-collision.cc:45-99) for ONE edge per CTA.
//
// Why: an edge that grazes an obstacle needs thousands of bisection samples, each depending on the previous one,
// and keeps tens of disjoint validated intervals on the binding body pairs.  The lane-per-edge walker pays the
// whole sample (FK, list scans, narrow phase) in one thread: ~110 us per sample on UR5, ~350 us on the 30-joint
// tree, and the longest edge is the floor of the batch time.  Here the 128 threads of a CTA share the sample:
//   A  joint transforms (thread per joint)  |  cover search: the validated interval of each body pair that
//      contains t (thread per body pair; lists are 64-entry arrays in global memory, L1-resident)
//   B  chain composition (one thread: 63 flops per joint)
//   C  geometry placements (thread per geometry)
//   D  narrow phase of the body pairs the reference certainly evaluates (no validated interval around t)
//   E  replay of validateIntervals in the reference order (one thread); a covered pair whose cover turns out
//      not to contain the running intersection needs a second narrow-phase stage (speculative: results of
//      pairs the replay then skips are discarded, never inserted)
//   F  interval inserts (thread per body pair), smallest-first swap, valid subset, next parameter.
// Results are bit-identical to the walker and the oracle: same per-pair arithmetic, same order of decisions.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>
#include <mutex>
#include <tuple>

#include "dev_intervals.cuh"

namespace hppb200 {

#ifndef HPPB200_DT_THREADS
#define HPPB200_DT_THREADS 128
#endif
constexpr int DT_THREADS = HPPB200_DT_THREADS;
#ifndef HPPB200_DT_MINB
#define HPPB200_DT_MINB 3
#endif
constexpr int DT_MINB = HPPB200_DT_MINB;   // CTAs per SM the register budget is held to
constexpr int DT_K = 64;     // validated intervals per body pair (lists in global memory)
constexpr int DT_KS = 256;   // intervals of the edge's valid subset (shared memory)
constexpr uint32_t DT_SMEM_LISTS = 64u * 1024u;  // shared-memory budget of the per-pair lists of the first pass

struct TailArgs {
  const double* q1;
  const double* q2;
  const double* L;
  const uint32_t* edge_list;  // optional indirection
  size_t n;
  int reverse;
  double tol;
  uint32_t max_iter;
  int overflow_reason;
  uint8_t* valid;
  double* t_fail;
  double* t_last;
  int32_t* pair;
  uint8_t* reason;
  unsigned long long* counter;
  double2* lists;  // [grid][nbodypairs][K] in global memory, or null: the lists live in shared memory
  int K;           // validated intervals per body pair
  double2* retry_lists;  // lists in shared memory with K < DT_K: [grid][nbodypairs][DT_K] in global memory for the edges that outgrow them
  unsigned long long* clk;  // optional [8]: cycles of thread 0 per phase (HPPB200_WF_TRACE)
  unsigned long long* cta_log;  // optional [grid][4] (HPPB200_WF_TRACE): exit time, start / duration / samples of the CTA's longest edge (globaltimer ns)
  TailResume resume;        // Progressive: continue the edges of these wavefront slots (slots == null: start edges afresh)
};

// dynamic shared memory of one CTA
struct TailLayout {
  uint32_t off_plc, off_li, off_sv, off_vmax, off_cover, off_res, off_pcov, off_pres, off_sub, off_lists, off_order, off_cnt, off_state,
      off_pflag, off_depth, off_visit, off_gdist, off_sorted, off_bodyof, off_jobs, total;
};
// smemK: entries per body pair of the lists kept in shared memory (0: lists in global memory)
__host__ __device__ inline TailLayout tailLayout(const ModelView& mv, int smemK) {
  TailLayout Lo;
  uint32_t o = 0;
  Lo.off_plc = o; o += (uint32_t)(mv.ngeoms + mv.njoints) * 12u * 8u;
  Lo.off_li = o; o += (uint32_t)mv.njoints * 12u * 8u;
  Lo.off_sv = o; o += (uint32_t)mv.nv * 8u;
  Lo.off_vmax = o; o += (uint32_t)mv.nbodypairs * 8u;
  o = (o + 15u) & ~15u;  // double2 records
  Lo.off_cover = o; o += (uint32_t)mv.nbodypairs * 16u;
  Lo.off_res = o; o += (uint32_t)mv.nbodypairs * 16u;
  Lo.off_pcov = o; o += (uint32_t)mv.nbodypairs * 16u;  // the same two records by order position (replay)
  Lo.off_pres = o; o += (uint32_t)mv.nbodypairs * 16u;
  Lo.off_sub = o; o += (uint32_t)DT_KS * 16u;
  Lo.off_lists = o; o += (uint32_t)mv.nbodypairs * (uint32_t)smemK * 16u;
  Lo.off_order = o; o += (uint32_t)mv.nbodypairs * 2u;
  Lo.off_cnt = o; o += (uint32_t)mv.nbodypairs * 2u;
  Lo.off_state = o; o += (uint32_t)mv.nbodypairs;
  Lo.off_pflag = o; o += (uint32_t)mv.nbodypairs + ((uint32_t)mv.nbodypairs + 31u) / 32u;  // + first valid_ lane per block
  Lo.off_depth = o; o += (uint32_t)mv.njoints;
  o = (o + 3u) & ~3u;
  Lo.off_visit = o; o += ((uint32_t)mv.nbodypairs + 31u) / 32u * 4u;  // positions the replay has to visit
  // narrow phase by GEOMETRY pair: distance per geometry pair, the static cost order, the owner body pair, this stage's jobs
  o = (o + 7u) & ~7u;
  Lo.off_gdist = o; o += (uint32_t)mv.nbgpairs * 8u;
  Lo.off_sorted = o; o += (uint32_t)mv.nbgpairs * 2u;
  Lo.off_bodyof = o; o += (uint32_t)mv.nbgpairs * 2u;
  Lo.off_jobs = o; o += (uint32_t)mv.nbgpairs * 2u;
  Lo.total = (o + 15u) & ~15u;
  return Lo;
}

enum : uint8_t { ST_VALID = 1, ST_CAND = 2, ST_DONE = 4, ST_INSERT = 8 };
enum : int { RUN_SAMPLE = 0, RUN_STAGE = 1, RUN_FINISHED = 2 };
enum : uint8_t { PF_RESULT = 1, PF_HIT = 2, PF_SKIP = 4, PF_FOLD = 8 };  // per order position, replay in / out

__device__ __forceinline__ void store12(double* p, const T3& X) {
#pragma unroll
  for (int i = 0; i < 9; ++i) p[i] = X.r[i];
  p[9] = X.t.x; p[10] = X.t.y; p[11] = X.t.z;
}

template <int METHOD, int THREADS = 128>
__global__ void __launch_bounds__(THREADS, THREADS == 96 ? 4 : hppb200::DT_MINB)
edgeTailKernel(ModelView mv, TailArgs A, DeviceStats* stats) {
  // THREADS = 96 (three warps, four CTAs per SM at the same 168 registers): Progressive on models with many body pairs, whose
  // two-entry lists leave the shared memory for a fourth CTA (63.6 -> 52.6 ms on 87 k edges of the 273-pair humanoid; Dichotomy
  // does not gain: its lists would have to shrink to 4 entries)
  constexpr int DT_THREADS = THREADS;
  constexpr bool PROG = METHOD == HPPB200_METHOD_PROGRESSIVE;
  extern __shared__ __align__(16) unsigned char smem[];
  const Tables T = makeTables(mv.blob, mv);  // global copy (read-only, L1 resident)
  const TailLayout Lo = tailLayout(mv, A.lists ? 0 : A.K);
  double* plc = (double*)(smem + Lo.off_plc);  // geometry g at 12 g, joint j (1-based) at 12 (ngeoms + j - 1)
  double* li = (double*)(smem + Lo.off_li);
  double* sv = (double*)(smem + Lo.off_sv);
  double* vmax = (double*)(smem + Lo.off_vmax);
  double2* cover = (double2*)(smem + Lo.off_cover);
  double2* res = (double2*)(smem + Lo.off_res);
  uint16_t* order = (uint16_t*)(smem + Lo.off_order);
  uint16_t* lcnt = (uint16_t*)(smem + Lo.off_cnt);
  uint8_t* state = (uint8_t*)(smem + Lo.off_state);
  double2* pcov = (double2*)(smem + Lo.off_pcov);
  double2* pres = (double2*)(smem + Lo.off_pres);
  uint8_t* pflag = (uint8_t*)(smem + Lo.off_pflag);  // by position: PF_* in, decision out
  uint8_t* jdepth = (uint8_t*)(smem + Lo.off_depth);
  uint32_t* visit = (uint32_t*)(smem + Lo.off_visit);
  double* gdist = (double*)(smem + Lo.off_gdist);
  uint16_t* sortedGp = (uint16_t*)(smem + Lo.off_sorted);
  uint16_t* bodyOf = (uint16_t*)(smem + Lo.off_bodyof);
  uint16_t* jobs = (uint16_t*)(smem + Lo.off_jobs);
  __shared__ int s_cls[4], s_njobs, s_firstLimit;
  __shared__ double s_t, s_ilo, s_ihi, s_last;
  __shared__ int s_run, s_pos, s_smallest, s_subCnt, s_overflow, s_has, s_stop, s_depth, s_finished;
  __shared__ uint32_t s_edge, s_iters, s_slot;
  __shared__ int s_redo;  // the current edge outgrew the shared-memory lists: the CTA validates it again with DT_K entries per pair in global memory

  const int tid = threadIdx.x;
  const int NP = mv.nbodypairs, NJ = mv.njoints, NG = mv.ngeoms;
  const double INF = CUDART_INF;
  const bool reverse = A.reverse != 0;
  int K = A.K;
  double2* const baseLists = A.lists ? A.lists + (size_t)blockIdx.x * NP * A.K : (double2*)(smem + Lo.off_lists);
  double2* lists = baseLists;
  double2* sub = (double2*)(smem + Lo.off_sub);
  __shared__ long long s_ck[8];
  long long c0 = 0;
  const bool timing = A.clk != nullptr && tid == 0;
  unsigned long long logStart = 0, logDur = 0, logIters = 0;
  if (timing)
    for (int i = 0; i < 8; ++i) s_ck[i] = 0;
#define DT_TICK(i) do { if (timing) { const long long c1 = clock64(); s_ck[i] += c1 - c0; c0 = c1; } } while (0)
  GjkCounters cnt = {0u, 0u, 0u};
  unsigned tests = 0, samples = 0, caps = 0, overflows = 0, discarded = 0;

  if (tid == 0) {  // tree level of every joint (parents precede children)
    int deepest = 0;
    for (int j = 1; j <= NJ; ++j) {
      const int parent = T.joints[j - 1].parent;
      const int d = parent == 0 ? 1 : jdepth[parent - 1] + 1;
      jdepth[j - 1] = (uint8_t)d;
      deepest = d > deepest ? d : deepest;
    }
    s_depth = deepest;
  }
  // Static cost order of the geometry pairs, once per CTA: GJK with two convex shapes, one, none, closed forms.  A stage's
  // jobs keep this order, so a round of 128 jobs is as uniform in cost (and in code path) as the stage allows: the
  // narrow phase takes sum over rounds of the SLOWEST job of the round (thread per body pair, in pair order, it was
  // ~3 x a convex-convex GJK per sample on the 30-joint model).
  const int NBG = mv.nbgpairs;
  auto gpClass = [&](int gp) -> int {
    const DevPair& pr = T.bgpairs[gp];
    if (pr.kind != PK_GJK) return 3;
    const DevShape& sa = *(const DevShape*)(pr.a >= 0 ? &T.geoms[pr.a].shape : &T.obstacles[-(pr.a + 1)].shape);
    const DevShape& sb = *(const DevShape*)(pr.b >= 0 ? &T.geoms[pr.b].shape : &T.obstacles[-(pr.b + 1)].shape);
    return 2 - (sa.type == SH_CONVEX ? 1 : 0) - (sb.type == SH_CONVEX ? 1 : 0);
  };
  if (tid < 4) s_cls[tid] = 0;
  __syncthreads();
  for (int gp = tid; gp < NBG; gp += DT_THREADS) atomicAdd(&s_cls[gpClass(gp)], 1);
  for (int p = tid; p < NP; p += DT_THREADS)
    for (int gp = T.bodypairs[p].pair_begin; gp < T.bodypairs[p].pair_end; ++gp) bodyOf[gp] = (uint16_t)p;
  __syncthreads();
  if (tid == 0) {
    int acc = 0;
    for (int c = 0; c < 4; ++c) { const int k = s_cls[c]; s_cls[c] = acc; acc += k; }
  }
  __syncthreads();
  for (int gp = tid; gp < NBG; gp += DT_THREADS) sortedGp[atomicAdd(&s_cls[gpClass(gp)], 1)] = (uint16_t)gp;
  // First sample of a fresh edge: no pair has a validated interval, so every pair is a candidate, and the reference stops at
  // the FIRST colliding pair of the construction order.  When the pairs make more jobs than one round of threads, the first
  // stage takes only the leading pairs that fill one round; the replay then either meets a collision among them (most
  // random configurations of a cluttered scene collide: the other two thirds of the distances are never needed) or asks
  // for the rest through the second-stage protocol below.
  if (tid == 0) {
    int lim = NP;
    if (NBG > DT_THREADS) {
      lim = 0;
      while (lim < NP && T.bodypairs[lim].pair_end <= DT_THREADS) ++lim;
      if (lim == 0) lim = NP;
    }
    s_firstLimit = lim;
    s_redo = 0;
  }
  __syncthreads();
  for (;;) {
    if (tid == 0) {
      if (s_redo) s_has = 1;  // the same edge (and slot) once more, from its start
      else {
        const unsigned long long i = atomicAdd(A.counter, 1ull);
        s_has = i < A.n ? 1 : 0;
        if (s_has) {
          if (A.resume.slots) { s_slot = A.resume.slots[i]; s_edge = A.resume.st_edge[s_slot]; }
          else s_edge = A.edge_list ? A.edge_list[i] : (uint32_t)i;
        }
      }
    }
    __syncthreads();
    if (!s_has) break;
    {
      // In-kernel second level: an edge that outgrows the K-entry lists in shared memory (K is cut to 8 on whole batches for a
      // third resident CTA) is validated again at once by the CTA that found out, with DT_K entries per pair in global memory.
      // (A second LAUNCH for those edges, 131 in a million on BASELINE configs[3], waited for the whole batch and then ran the
      // serial chain of the longest of them alone: 23 ms behind 53 ms.)  s_redo is cleared behind the next barriers.
      const bool redo = s_redo != 0;
      K = redo ? DT_K : A.K;
      lists = redo ? A.retry_lists + (size_t)blockIdx.x * NP * DT_K : baseLists;
    }
    const uint32_t edge = s_edge;
    unsigned long long tEdge0 = 0;
    if (timing && A.cta_log) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tEdge0));
    const double L = A.L[edge];
    const double* q1 = A.q1 + (size_t)edge * mv.nq;
    const double* q2 = A.q2 + (size_t)edge * mv.nq;
    // dichotomy.cc:113-121: the valid part ends where the first (last, if reverse) validated interval ends
    // Progressive: lastValidTime = the last sampled valid parameter (progressive.cc:113-114)
    auto dichLast = [&]() -> double { return PROG ? s_last : (reverse ? sub[s_subCnt - 1].x : sub[0].y); };
    auto finishEdge = [&](bool ok, int rs, double tf, double tl, int pr) {
      A.valid[edge] = ok ? 1 : 0;
      if (A.t_fail) A.t_fail[edge] = tf;
      if (A.t_last) A.t_last[edge] = tl;
      if (A.pair) A.pair[edge] = pr;
      if (A.reason) A.reason[edge] = (uint8_t)rs;
    };

    // ---- BodyPairCollision::setupPath for every body pair (body-pair-collision.cc:101-117) ----
    for (int p = tid; p < NP; p += DT_THREADS) { order[p] = (uint16_t)p; lcnt[p] = 0; state[p] = 0; }
    if (L != 0)
      for (int j = tid; j < NJ; j += DT_THREADS) {
        const DevJoint& J = T.joints[j];
        sv[J.idx_v] = jointVelocityNorm(q1, q2, J.idx_q, J.type == HPPB200_JOINT_FREEFLYER ? 6 : 1, L);
      }
    __syncthreads();
    for (int p = tid; p < NP; p += DT_THREADS) {  // SolidSolidCollision::computeMaximalVelocity (:76-99)
      const DevBodyPair& B = T.bodypairs[p];
      double Vm;
      if (L == 0) Vm = INF;
      else {
        double ab = 0;
        for (int c = B.coef_begin; c < B.coef_end; ++c) ab += T.coefs[c].value * sv[T.coefs[c].idx_v];
        Vm = ab;
        if (B.joint_b != 0) {
          double ba = 0;
          for (int c = B.rcoef_begin; c < B.rcoef_end; ++c) ba += T.rcoefs[c].value * sv[T.rcoefs[c].idx_v];
          if (ba < ab) Vm = ba;
        }
      }
      vmax[p] = Vm;
    }
    if (tid == 0) {
      const double t = reverse ? L : 0.0;
      s_t = t;
      sub[0] = make_double2(t, t);  // validSubset.unionInterval((t, t)) (dichotomy.cc:84)
      s_subCnt = 1;
      s_iters = 0;
      s_run = RUN_SAMPLE;
      s_last = t;                                            // progressive.cc:93-100
      s_finished = (reverse ? t <= 0.0 : t >= L) ? 1 : 0;  // a zero-length path: one configuration to validate
      s_redo = 0;
    }
    if (PROG && A.resume.slots) {
      // continue an edge of the wavefront engine (kernels_wavefront.cu) at a sample boundary: order, one validated
      // interval per pair (by order position there, by pair here), valid_ bits, t, lastValidTime, counters
      __syncthreads();
      const TailResume& R = A.resume;
      const uint32_t e = s_slot;
      for (int pos = tid; pos < NP; pos += DT_THREADS) {
        const int id = R.ord[(size_t)e * R.ordStride + pos];
        order[pos] = (uint16_t)id;
        const double2 lh = R.lohi[(size_t)pos * R.E + e];
        const bool isValid = (R.st_vmask[(size_t)(pos >> 6) * R.E + e] >> (pos & 63)) & 1ull;
        lists[(size_t)id * K] = lh;
        lcnt[id] = (lh.x > lh.y) ? 0 : 1;
        state[id] = isValid ? ST_VALID : 0;
      }
      if (tid == 0) {
        s_t = R.st_t[e];
        s_last = R.st_last[e];
        s_iters = R.st_iters[e];
        s_finished = (R.st_flags[e] >> 2) & 3;
      }
    }
    __syncthreads();

    for (;;) {  // one sample per turn
      if (tid == 0) {
        if (s_iters >= A.max_iter) {
          ++caps;
          finishEdge(false, HPPB200_REASON_ITER_CAP, s_t, dichLast(), -1);
          s_run = RUN_FINISHED;
        } else {
          ++s_iters;
          ++samples;
        }
      }
      __syncthreads();
      if (s_run == RUN_FINISHED) break;
      const double t = s_t;
      if (timing) c0 = clock64();

      // ---- A: joint transforms | cover search ----
      {
        const EdgeQ eq = edgeQ(q1, q2, t, L);
        for (int j = tid; j < NJ; j += DT_THREADS) store12(li + j * 12, jointLocal(T.joints[j], eq));
      }
      for (int p = tid; p < NP; p += DT_THREADS) {
        uint8_t st = state[p] & ST_VALID;
        if (!st) {
          const int c = lcnt[p];
          const double2* lst = lists + (size_t)p * K;
          double2 cv = make_double2(INF, -INF);
          if (!reverse) {  // t sits right behind the validated start: the interval around it is near the front
            for (int i = 0; i < c; ++i) {
              const double2 v = lst[i];
              if (v.y >= t) { if (v.x <= t) cv = v; break; }
            }
          } else {
            for (int i = c - 1; i >= 0; --i) {
              const double2 v = lst[i];
              if (v.x <= t) { if (t <= v.y) cv = v; break; }
            }
          }
          cover[p] = cv;
          if (cv.y == -INF) st = ST_CAND;  // no validated interval around t: evaluated whatever the order
        }
        state[p] = st;
      }
      __syncthreads();
      DT_TICK(0);
      // ---- B: oMi = oMi[parent] * liMi, one tree level at a time (a lane per joint of the level) ----
      if (tid < 32) {
        if (s_depth == NJ) {  // a chain: nothing to share out
          if (tid == 0)
            for (int j = 1; j <= NJ; ++j) {
              const T3 l = load12(li + (j - 1) * 12);
              store12(plc + (NG + j - 1) * 12, j == 1 ? l : compose(load12(plc + (NG + j - 2) * 12), l));
            }
        } else
        for (int lv = 1; lv <= s_depth; ++lv) {
          for (int j = tid + 1; j <= NJ; j += 32) {
            if (jdepth[j - 1] != lv) continue;
            const int parent = T.joints[j - 1].parent;
            const T3 l = load12(li + (j - 1) * 12);
            store12(plc + (NG + j - 1) * 12, parent == 0 ? l : compose(load12(plc + (NG + parent - 1) * 12), l));
          }
          __syncwarp();
        }
        if (tid == 0) { s_pos = 0; s_ilo = -INF; s_ihi = INF; s_smallest = 0; s_overflow = 0; s_stop = 0; }
      }
      __syncthreads();
      DT_TICK(1);
      // ---- C: oMg = oMi[joint] * placement ----
      for (int g = tid; g < NG; g += DT_THREADS) {
        const DevGeom& G = T.geoms[g];
        store12(plc + g * 12, G.joint == 0 ? load12(G.P) : compose(load12(plc + (NG + G.joint - 1) * 12), load12(G.P)));
      }
      __syncthreads();
      DT_TICK(2);

      // ---- D / E: narrow phase of the flagged pairs, then the replay; a second stage when the replay meets a
      //      covered pair that the reference evaluates after all ----
      const bool freshFirst = s_iters == 1 && !(PROG && A.resume.slots) && s_firstLimit < NP;  // (order = construction order)
      for (int stage = 0;; ++stage) {
        const int plim = (freshFirst && stage == 0) ? s_firstLimit : NP;  // body pairs (= positions) this stage evaluates
        // D1: this stage's jobs = the geometry pairs of the flagged body pairs, in the static cost order
        if (tid == 0) s_njobs = 0;
        __syncthreads();
        for (int base = 0; base < NBG; base += DT_THREADS) {
          const int i = base + tid;
          int gp = 0;
          bool want = false;
          if (i < NBG) {
            gp = sortedGp[i];
            want = (int)bodyOf[gp] < plim && (state[bodyOf[gp]] & (ST_CAND | ST_DONE)) == ST_CAND;
          }
          const unsigned m = __ballot_sync(0xffffffffu, want);
          int off = 0;
          if ((tid & 31) == 0 && m) off = atomicAdd(&s_njobs, __popc(m));
          off = __shfl_sync(0xffffffffu, off, 0);
          if (want) jobs[off + __popc(m & ((1u << (tid & 31)) - 1u))] = (uint16_t)gp;
        }
        __syncthreads();
        // D2: one distance per thread and round.  (The reference stops a body pair at its first colliding geometry pair;
        // the later distances computed here are never read: D3 walks them in order.)
        {
          const int nj = s_njobs;
          for (int j = tid; j < nj; j += DT_THREADS) {
            const int gp = jobs[j];
            gdist[gp] = pairLowerBound(T, T.bgpairs[gp], plc, 1, mv.vertices, cnt);
          }
        }
        __syncthreads();
        // D3: computeDistanceLowerBound per body pair (:175-198): geometry pairs in order, the first collision reports
        for (int p = tid; p < plim; p += DT_THREADS) {
          if ((state[p] & (ST_CAND | ST_DONE)) != ST_CAND) continue;
          const DevBodyPair& B = T.bodypairs[p];
          double dlb = INF;
          int hit = -1;
          for (int gp = B.pair_begin; gp < B.pair_end; ++gp) {
            const double d = gdist[gp];
            ++tests;
            if (d <= HPPB200_COLLISION_THRESHOLD) { hit = gp; break; }
            if (d < dlb) dlb = d;
          }
          if (hit >= 0) res[p] = make_double2((double)hit, -INF);
          else {
            double hd, ht;  // :72-85
            if (dlb == INF) { hd = INF; ht = 0; }
            else {
              const double Vm = vmax[p];
              hd = dlb / Vm;
              ht = (Vm != 0) ? 2 * A.tol / Vm : INF;
            }
            res[p] = make_double2(t - (hd + ht), t + (hd + ht));
          }
          state[p] |= ST_DONE;
        }
        __syncthreads();
        DT_TICK(3);
        // validateIntervals in the current order.  (1) every thread lays the records out by order position;
        // (2) one thread walks them: the running intersection is a serial dependence, so the walk is kept to a
        // few selects per position (a valid_ pair is a pair with the result [0, L] and no cover);
        // (3) the decisions are applied position-parallel below.
        // Only the pairs that are not valid_ need a visit, plus the first valid_ one (it clips the running
        // interval to [0, L]; the later ones change nothing).
        for (int base = (s_pos & ~31) + (tid & ~31); base < NP; base += DT_THREADS) {
          const int pos = base + (tid & 31);
          bool nonValid = false, isValid = false;
          if (pos >= s_pos && pos < NP) {
            const int p = order[pos];
            const uint8_t st = state[p];
            double2 c = make_double2(INF, -INF), r = make_double2(0.0, L);
            uint8_t f = PF_RESULT;
            if (!(st & ST_VALID)) {
              c = cover[p];
              f = 0;
              if (st & ST_DONE) { r = res[p]; f = (r.y == -INF) ? (PF_RESULT | PF_HIT) : PF_RESULT; }
              nonValid = true;
            } else isValid = true;
            pcov[pos] = c; pres[pos] = r; pflag[pos] = f;
          }
          const unsigned nv = __ballot_sync(0xffffffffu, nonValid), vv = __ballot_sync(0xffffffffu, isValid);
          if ((tid & 31) == 0) { visit[base >> 5] = nv; pflag[NP + (base >> 5)] = vv ? (uint8_t)(__ffs(vv)) : (uint8_t)0; }
        }
        __syncthreads();
        if (tid == 0) {
          int pos = s_pos, smallest = s_smallest, run = RUN_SAMPLE;
          double ilo = s_ilo, ihi = s_ihi;
          bool clipped = ilo >= 0.0 && ihi <= L;  // a valid_ pair was folded already
          const int words = (NP + 31) >> 5;
          int stop = NP;
          for (int w = pos >> 5; w < words && run == RUN_SAMPLE; ++w) {
            uint32_t m = visit[w];
            const int fv = pflag[NP + w];  // 1 + lane of the first valid_ pair of this block, 0: none
            if (fv && !clipped) { m |= 1u << (fv - 1); clipped = true; }
            while (m) {
              const int b = __ffs(m) - 1;
              m &= m - 1u;
              pos = (w << 5) + b;
              const double2 c = pcov[pos], r = pres[pos];
              const uint8_t f = pflag[pos];
              const bool skip = c.x <= ilo && ihi <= c.y;  // icl::contains(validInterval_, running interval) (:56-63)
              if (!skip && f != PF_RESULT) {  // evaluated by the reference: not computed yet, or in collision
                if (f & PF_HIT) { finishEdge(false, HPPB200_REASON_COLLISION, t, dichLast(), (int)r.x); run = RUN_FINISHED; }
                else run = RUN_STAGE;
                stop = pos;
                break;
              }
              pflag[pos] = skip ? PF_SKIP : PF_FOLD;
              smallest = (!skip && ihi > r.y) ? pos : smallest;  // continuous-validation.hh:283
              ilo = (!skip && ilo < r.x) ? r.x : ilo;
              ihi = (!skip && r.y < ihi) ? r.y : ihi;
            }
          }
          pos = stop;
          s_stop = pos; s_smallest = smallest; s_ilo = ilo; s_ihi = ihi; s_run = run;
        }
        __syncthreads();
        if (s_run == RUN_STAGE) {
          // flag the pair the walk stopped at and every later pair that can still be evaluated (the running
          // interval only shrinks, so a cover that contains it now is final); positions before it are decided
          const double ilo = s_ilo, ihi = s_ihi;
          for (int pos = s_stop + tid; pos < NP; pos += DT_THREADS) {
            const int p2 = order[pos];
            const uint8_t s2 = state[p2];
            if (s2 & (ST_VALID | ST_DONE)) continue;
            const double2 c2 = cover[p2];
            if (!(c2.x <= ilo && ihi <= c2.y)) state[p2] = s2 | ST_CAND;
          }
        }
        // decisions of the positions walked in this turn
        for (int pos = s_pos + tid; pos < s_stop; pos += DT_THREADS) {
          const int p = order[pos];
          const uint8_t st = state[p];
          if (st & ST_VALID) continue;
          if (pflag[pos] == PF_SKIP) { if (st & ST_DONE) ++discarded; state[p] = 0; }
          else state[p] = st | ST_INSERT;
        }
        __syncthreads();
        if (tid == 0) s_pos = s_stop;
        __syncthreads();
        DT_TICK(4);
        if (s_run != RUN_STAGE) break;
        if (timing) ++s_ck[7];
      }
      if (s_run == RUN_FINISHED) break;

      // ---- F: validInterval_.insert (:90), valid_ (:92-96) ----
      for (int p = tid; p < NP; p += DT_THREADS) {
        if (!(state[p] & ST_INSERT)) continue;
        int c = lcnt[p];
        double2* lst = lists + (size_t)p * K;
        const double2 r = res[p];
        if (!wdUnion(lst, 1, c, K, r.x, r.y)) { s_overflow = 1; continue; }
        lcnt[p] = (uint16_t)c;
        state[p] = wdContains(lst, 1, c, 0.0, L) ? ST_VALID : 0;
      }
      __syncthreads();
      DT_TICK(5);
      if (tid == 0) {
        const double ilo = s_ilo, ihi = s_ihi;
        int sc = s_subCnt;
        bool over = s_overflow != 0;
        if (!over) {
          // src/continuous-validation.cc:144-146: smallest interval first
          const int smallest = s_smallest;
          if (NP > 1 && smallest != 0) { const uint16_t a0 = order[0]; order[0] = order[smallest]; order[smallest] = a0; }
          if (!PROG) over = !wdUnion(sub, 1, sc, DT_KS, ilo, ihi);  // dichotomy.cc:96-109
        }
        if (over) {
          if (s_overflow != 0 && A.retry_lists && K != DT_K) s_redo = 1;  // (a pair list, not the valid subset)
          else {
            if (A.overflow_reason == HPPB200_REASON_ITER_CAP) ++caps; else ++overflows;
            finishEdge(false, A.overflow_reason, t, dichLast(), -1);
          }
          s_run = RUN_FINISHED;
        } else if (PROG) {  // progressive.cc:104-117
          double tn = reverse ? ilo : ihi;
          s_last = t;
          if (reverse ? tn <= 0.0 : tn >= L) { tn = reverse ? 0.0 : L; s_finished++; }
          if (s_finished >= 2) {
            finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
            s_run = RUN_FINISHED;
          } else {
            s_t = tn;
            s_run = RUN_SAMPLE;
          }
        } else {
          s_subCnt = sc;
          if (wdContains(sub, 1, sc, 0.0, L)) {
            finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
            s_run = RUN_FINISHED;
          } else {
            double t0, t1;  // middle of the first non-validated gap
            if (!reverse) { t0 = sub[0].y; t1 = sc > 1 ? sub[1].x : L; }
            else { t0 = sub[sc - 1].x; t1 = sc > 1 ? sub[sc - 2].y : 0.0; }
            s_t = .5 * (t0 + t1);
            s_run = RUN_SAMPLE;
          }
        }
      }
      __syncthreads();
      DT_TICK(6);
      if (s_run == RUN_FINISHED) break;
    }
    if (timing && A.cta_log) {
      unsigned long long t1;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
      if (t1 - tEdge0 > logDur) { logDur = t1 - tEdge0; logStart = tEdge0; logIters = s_iters; }
    }
    __syncthreads();  // the edge's shared state is dead from here
  }
  if (timing && A.cta_log) {
    unsigned long long t1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
    unsigned long long* o = A.cta_log + 4 * (size_t)blockIdx.x;
    o[0] = t1; o[1] = logStart; o[2] = logDur; o[3] = logIters;
  }

  if (timing)
    for (int i = 0; i < 8; ++i) atomicAdd(A.clk + i, (unsigned long long)s_ck[i]);
#undef DT_TICK
  if (stats) {
    if (tests) atomicAdd(&stats->pair_tests, (unsigned long long)tests);
    if (cnt.iterations) atomicAdd(&stats->gjk_iterations, (unsigned long long)cnt.iterations);
    if (cnt.cap_hits) atomicAdd(&stats->gjk_cap_hits, (unsigned long long)cnt.cap_hits);
    if (samples) atomicAdd(&stats->edge_samples, (unsigned long long)samples);
    if (caps) atomicAdd(&stats->iter_cap_hits, (unsigned long long)caps);
    if (overflows) atomicAdd(&stats->overflow, (unsigned long long)overflows);
    if (discarded) atomicAdd(&stats->discarded_tests, (unsigned long long)discarded);
  }
}

// Runs Dichotomy or Progressive on the edges of `d_list` (or on all n edges when d_list is null), one CTA per edge.  An edge whose
// lists outgrow the capacity is reported with `overflow_reason` (counted in DeviceStats::overflow).  smemLists:
// per-pair lists in shared memory (as many entries as DT_SMEM_LISTS holds, at most DT_K); else DT_K entries in global memory.
bool launchEdgeTail(Workspace* ws, int method, double tol, int reverse, const double* d_q1, const double* d_q2, const double* d_L,
                    const uint32_t* d_list, size_t n, uint8_t* d_valid, double* d_tfail, double* d_tlast, int32_t* d_pair,
                    uint8_t* d_reason, int overflow_reason, bool smemLists, const TailResume* resume) {
  const ModelView& mv = ws->view;
  // (see the THREADS note at the kernel; HPPB200_TAIL_THREADS=128 keeps four warps everywhere)
  bool narrowCta = method == HPPB200_METHOD_PROGRESSIVE && smemLists && mv.nbodypairs > 128 && n >= (size_t)ws->num_sms * 8;
  if (const char* v = std::getenv("HPPB200_TAIL_THREADS")) narrowCta = narrowCta && std::atoi(v) == 96;
  const int threads = narrowCta ? 96 : DT_THREADS, minb = narrowCta ? 4 : DT_MINB;
  auto kern = method == HPPB200_METHOD_PROGRESSIVE
                  ? (narrowCta ? edgeTailKernel<HPPB200_METHOD_PROGRESSIVE, 96> : edgeTailKernel<HPPB200_METHOD_PROGRESSIVE, 128>)
                  : edgeTailKernel<HPPB200_METHOD_DICHOTOMY, 128>;
  if (n == 0) return true;
  if (mv.nbodypairs > 65535 || mv.nbodypairs == 0) return false;
  if (mv.nbgpairs > 65535) return false;  // (16-bit job records)
  int K = DT_K;
  if (smemLists) {
    K = std::min<int>(DT_K, (int)(DT_SMEM_LISTS / (16u * (uint32_t)mv.nbodypairs)));
    if (K < 8) return false;
    // Progressive keeps ONE connected validated interval per body pair: two entries, the rest of the shared memory
    // goes to the L1 (an edge that ever needed more would be redone with DT_K entries like any other overflow)
    if (method == HPPB200_METHOD_PROGRESSIVE) K = std::min(K, 2);
    if (const char* v = std::getenv("HPPB200_TAIL_K")) K = std::max(2, std::min(K, std::atoi(v)));  // (tests: short lists, so that edges outgrow them)
  }
  TailLayout Lo = tailLayout(mv, smemLists ? K : 0);
  if (Lo.total > 200u * 1024u) return false;
  // (attribute and occupancy are host calls of several microseconds each: asked once per (device, method, shared-memory size);
  //  the one-item calls of a planner come through here)
  int perSM = 0;
  {
    static std::mutex mu;
    static std::map<std::tuple<int, int, int, uint32_t>, int> cache;
    std::lock_guard<std::mutex> lock(mu);
    const auto key = std::make_tuple(ws->device, method, threads, Lo.total);
    auto it = cache.find(key);
    if (it == cache.end()) {
      HPPB200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      HPPB200_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kern, threads, Lo.total));
      it = cache.emplace(key, perSM).first;
    }
    perSM = it->second;
  }
  if (perSM < 1) return false;
  // Many edges (a whole batch, not a handful of long ones): the kernel is latency-bound at 4 warps per CTA, so resident CTAs
  // are worth more than list capacity.  Take the shortest lists (>= 8 entries) that let one more CTA onto the SM; an edge
  // that outgrows them is reported with `overflow_reason` and redone by the next level, like any other overflow.
  int tailOcc = 1;
  if (const char* v = std::getenv("HPPB200_TAIL_OCC")) tailOcc = std::atoi(v);
  if (tailOcc && smemLists && K > 8 && n >= (size_t)ws->num_sms * 8 && perSM < minb) {
    for (int k2 = K - 1; k2 >= 8; --k2) {
      const TailLayout L2 = tailLayout(mv, k2);
      int p2 = 0;
      HPPB200_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p2, kern, threads, L2.total));
      if (p2 > perSM) {
        // ... and two entries less than that: what the lists give up goes to the L1 the three CTAs share (the tables are read
        // from global memory; BASELINE configs[3]: 8 entries 51.0 ms, 6 entries 48.5 ms, 4 entries 50.2 ms with more edges redone)
        K = std::max(6, k2 - 2); Lo = tailLayout(mv, K);
        HPPB200_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kern, threads, Lo.total));
        break;
      }
    }
  }
  const size_t grid = std::min<size_t>(n, (size_t)ws->num_sms * (size_t)perSM);
  const bool trace0 = std::getenv("HPPB200_WF_TRACE") != nullptr;
  const bool retry = smemLists && K < DT_K;
  const size_t listBytes = smemLists && !retry ? 0 : grid * (size_t)mv.nbodypairs * DT_K * sizeof(double2);
  ws->d_tail.reserve(listBytes + 128 + (trace0 ? grid * 32 : 0));
  const bool trace = trace0;
  TailArgs A{};
  A.q1 = d_q1; A.q2 = d_q2; A.L = d_L; A.edge_list = d_list; A.n = n;
  A.reverse = reverse; A.tol = tol; A.max_iter = ws->max_iterations; A.overflow_reason = overflow_reason;
  A.valid = d_valid; A.t_fail = d_tfail; A.t_last = d_tlast; A.pair = d_pair; A.reason = d_reason;
  A.lists = smemLists ? nullptr : (double2*)ws->d_tail.p;
  A.retry_lists = retry ? (double2*)ws->d_tail.p : nullptr;
  A.K = K;
  if (resume) A.resume = *resume;
  A.counter = (unsigned long long*)(ws->d_tail.p + listBytes);
  A.clk = trace ? A.counter + 1 : nullptr;
  A.cta_log = trace ? A.counter + 16 : nullptr;
  HPPB200_CUDA_CHECK(cudaMemsetAsync(A.counter, 0, 9 * sizeof(unsigned long long), ws->stream));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (trace) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, ws->stream); }
  kern<<<(unsigned)grid, threads, Lo.total, ws->stream>>>(mv, A, ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  if (trace) {
    cudaEventRecord(e1, ws->stream);
    unsigned long long ck[8];
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ck, A.clk, sizeof(ck), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    float kms = 0;
    cudaEventElapsedTime(&kms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    std::fprintf(stderr, "[tail] kernel %.3f ms\n", kms);
    if (grid > 1) {
      std::vector<unsigned long long> lg(4 * grid);
      HPPB200_CUDA_CHECK(cudaMemcpy(lg.data(), A.cta_log, lg.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
      unsigned long long t0 = ~0ull, tEnd = 0;
      for (size_t b = 0; b < grid; ++b) { if (lg[4 * b + 1]) t0 = std::min(t0, lg[4 * b + 1]); tEnd = std::max(tEnd, lg[4 * b]); }
      std::vector<double> ends;
      for (size_t b = 0; b < grid; ++b) ends.push_back((double)(lg[4 * b] - t0) * 1e-6);
      std::vector<size_t> idx(grid);
      for (size_t b = 0; b < grid; ++b) idx[b] = b;
      std::sort(idx.begin(), idx.end(), [&](size_t a, size_t b) { return ends[a] > ends[b]; });
      std::vector<double> se = ends;
      std::sort(se.begin(), se.end());
      std::fprintf(stderr, "[tail] CTA exit times (ms after the first edge started): median %.2f, p90 %.2f, p99 %.2f, max %.2f; the last CTAs:\n", se[grid / 2],
                   se[grid * 9 / 10], se[grid * 99 / 100], se[grid - 1]);
      for (size_t k = 0; k < std::min<size_t>(6, grid); ++k) {
        const size_t b = idx[k];
        std::fprintf(stderr, "[tail]   CTA %zu exits at %.2f ms; its longest edge: started %.2f ms, took %.2f ms, %llu samples\n", b, ends[b],
                     (double)(lg[4 * b + 1] - t0) * 1e-6, (double)lg[4 * b + 2] * 1e-6, lg[4 * b + 3]);
      }
    }
    std::fprintf(stderr, "[tail] n=%zu grid=%zu (%d per SM) K=%d %s smem=%u B; thread-0 Mcycles: A %.2f B %.2f C %.2f narrow %.2f replay %.2f insert %.2f next %.2f; second stages %llu\n",
                 n, grid, perSM, K, smemLists ? "shared" : "global", Lo.total, ck[0] * 1e-6, ck[1] * 1e-6, ck[2] * 1e-6, ck[3] * 1e-6, ck[4] * 1e-6,
                 ck[5] * 1e-6, ck[6] * 1e-6, ck[7]);
  }
  return true;
}

}  // namespace hppb200

namespace hppb200 {
bool launchDichotomyTail(Workspace* ws, double tol, int reverse, const double* d_q1, const double* d_q2, const double* d_L,
                         const uint32_t* d_list, size_t n, uint8_t* d_valid, double* d_tfail, double* d_tlast, int32_t* d_pair,
                         uint8_t* d_reason, int overflow_reason, bool smemLists) {
  return launchEdgeTail(ws, HPPB200_METHOD_DICHOTOMY, tol, reverse, d_q1, d_q2, d_L, d_list, n, d_valid, d_tfail, d_tlast, d_pair, d_reason,
                        overflow_reason, smemLists, nullptr);
}
}  // namespace hppb200
