// Continuous validation engine for continuousValidation::Progressive
// (reference src/continuous-validation/progressive.cc:78-129) over many StraightPath edges:
// the production path when the per-edge state fits in shared memory (UR5-class robots);
// kernels_edges.cu keeps the generic lane-per-edge walker for everything else.
//
// Why a second design: in the walker every lane runs its own edge's sequential loop, so a warp
// executes FK, interval bookkeeping, closed forms and GJK of 32 unrelated edges interleaved
// (measured: 5 of 32 lanes active on average, 4 warps per SM).  Here the two kinds of work are
// separated and each runs in full warps:
//
//   control phase (warp 0 of the CTA, lane = edge slot): replays the reference's loop over the
//     body pairs of the current sample -- ContinuousValidation::validateConfiguration
//     (src/continuous-validation.cc:128-148), validateIntervals
//     (include/hpp/core/continuous-validation.hh:267-298), BodyPairCollision::validateConfiguration
//     (src/continuous-validation/body-pair-collision.cc:45-99) -- consuming narrow-phase results
//     that are already there, then advances t (progressive.cc:104-117), runs FK for the next
//     sample (uniform across the warp) and queues the narrow-phase tests the next replay will need;
//   narrow phase (all warps of the CTA): one geometry pair per thread from a shared-memory job
//     queue, GJK jobs and closed-form jobs in separate ends of the queue so that warps are uniform.
//
// Which tests to queue is a PREDICTION; what the results mean is decided by the replay, which
// follows the reference order exactly: a body pair the reference would skip (valid_ or validated
// interval containing the running intersection, body-pair-collision.cc:51-63) is skipped even if
// its distance was computed (the result is discarded), and a pair the reference evaluates whose
// distance is not there yet stops the replay until the next round.  The distance of (edge, t, pair)
// is a pure function, so flags, parameters and interval sets are bit-identical to the sequential
// algorithm whatever was predicted.  Candidates: every pair whose validated interval does not
// contain [t - hp, t + hp] intersected with the running interval, hp = alpha x the previous step.
//
// State: interval sets (one closed interval per body pair: Progressive keeps the validated set
// connected), pair order, valid_ bits, per-joint velocity norms and the result
// slots live in shared memory, SoA over the 32 slots; geometry placements go to an L2-resident
// global scratch, AoS per slot (written once per sample by the control lane, read 96 B at a time
// by the jobs).  Several CTAs per SM run out of phase, which overlaps control and narrow phase.
#include "dev_model.cuh"

namespace hppb200 {

constexpr int PE_SLOTS = 32;                   // edge slots per CTA = lanes of the control warp
constexpr int PE_RMAX = 12;                    // body-pair results per slot per round
constexpr int PE_QCAP = PE_SLOTS * PE_RMAX;    // job queue capacity

struct PEArgs {
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
  unsigned long long* counter;  // next edge to hand out
  double* scratch;              // [gridDim.x * PE_SLOTS][fkSlots * 12]
  int fkSlots;
  int overflow_reason;
  double alpha;
};

struct PELayout {
  uint32_t off_blob, off_lo, off_hi, off_S, off_resA, off_resB, off_t, off_valid, off_queue, off_order, off_emPos, total;
  int validWords;
};

struct PECtl {
  int nG, nC, head, done;
};

static PELayout peLayout(const ModelView& mv) {
  PELayout Lo{};
  uint32_t o = 32;  // [0,8) mbarrier, [16,32) PECtl
  Lo.off_blob = o; o += mv.edge_blob_bytes;
  o = (o + 15u) & ~15u;
  Lo.off_lo = o; o += (uint32_t)mv.nbodypairs * PE_SLOTS * 8u;
  Lo.off_hi = o; o += (uint32_t)mv.nbodypairs * PE_SLOTS * 8u;
  Lo.off_S = o; o += (uint32_t)mv.nv * PE_SLOTS * 8u;
  Lo.off_resA = o; o += PE_RMAX * PE_SLOTS * 8u;
  Lo.off_resB = o; o += PE_RMAX * PE_SLOTS * 8u;
  Lo.off_t = o; o += 2 * PE_SLOTS * 8u;  // t, L per slot (read by the jobs)
  Lo.validWords = (mv.nbodypairs + 31) / 32;
  Lo.off_valid = o; o += (uint32_t)Lo.validWords * PE_SLOTS * 4u;
  Lo.off_queue = o; o += PE_QCAP * 4u;
  Lo.off_order = o; o += (uint32_t)mv.nbodypairs * PE_SLOTS;
  Lo.off_emPos = o; o += PE_RMAX * PE_SLOTS;
  Lo.total = (o + 15u) & ~15u;
  return Lo;
}

template <bool TREE, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
progressiveEngineKernel(ModelView mv, PEArgs A, PELayout SL, DeviceStats* stats) {
  extern __shared__ __align__(16) unsigned char smem[];  // @region setup
  uint64_t* bar = (uint64_t*)smem;
  PECtl* ctl = (PECtl*)(smem + 16);
  unsigned char* sblob = smem + SL.off_blob;
  stageBlobTMA(sblob, mv.blob, mv.edge_blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);

  const unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int NP = mv.nbodypairs;
  const double INF = CUDART_INF;
  const bool reverse = A.reverse != 0;

  double* sLo = (double*)(smem + SL.off_lo) + lane;
  double* sHi = (double*)(smem + SL.off_hi) + lane;
  double* sSAll = (double*)(smem + SL.off_S);
  double* sS = sSAll + lane;
  double* sResAAll = (double*)(smem + SL.off_resA);
  double* sResBAll = (double*)(smem + SL.off_resB);
  double* sTAll = (double*)(smem + SL.off_t);
  uint32_t* sValid = (uint32_t*)(smem + SL.off_valid) + lane;
  uint32_t* queue = (uint32_t*)(smem + SL.off_queue);
  uint8_t* sOrder = smem + SL.off_order + lane;
  uint8_t* sEmPos = smem + SL.off_emPos + lane;
  const size_t slotStride = (size_t)A.fkSlots * 12;
  double* scratchCTA = A.scratch + (size_t)blockIdx.x * PE_SLOTS * slotStride;

  if (tid == 0) { ctl->nG = 0; ctl->nC = 0; ctl->head = 0; ctl->done = 0; }
  __syncthreads();

  // ---- control-lane state (warp 0 only) ---------------------------------------------------------
  bool active = false, retired = false, sampleOpen = false, fkDone = false;
  size_t e = 0;
  const double* q1 = nullptr;
  const double* q2 = nullptr;
  double L = 0, t = 0, lastValid = 0, ilo = -INF, ihi = INF, hp = INF, u = 0;
  int qmode = 0, pos = 0, smallest = 0, finished = 0, nEm = 0;
  unsigned iters = 0;
  unsigned samples = 0, caps = 0, overflows = 0, discarded = 0, ticks = 0;
  // narrow-phase counters (all threads)
  GjkCounters cnt = {0u, 0u, 0u};
  unsigned tests = 0;

  auto validBit = [&](int id) -> bool { return (sValid[(id >> 5) * PE_SLOTS] >> (id & 31)) & 1u; };
  auto finishEdge = [&](bool ok, int rs, double tf, double tl, int pr) {
    A.valid[e] = ok ? 1 : 0;
    if (A.t_fail) A.t_fail[e] = tf;
    if (A.t_last) A.t_last[e] = tl;
    if (A.pair) A.pair[e] = pr;
    if (A.reason) A.reason[e] = (uint8_t)rs;
    active = false;
    sampleOpen = false;
  };

  // One pass of the control warp over the pair positions, all lanes in step (one position per trip, trip
  // count decided by a vote, so the warp stays converged).  A lane starts in REPLAY mode: it follows
  // validateIntervals from `pos` on with the results queued last round (result k belongs to the k-th
  // queued position).  At the first pair the reference evaluates whose result is not there, the lane
  // switches to EMIT mode: that pair and the predicted candidates after it are recorded for the next
  // narrow phase.  A lane that reaches the end in REPLAY mode has completed the sample.
  int nGl = 0, nCl = 0;
  unsigned gmask = 0;
  auto scanPass = [&](bool on) {  // @region replay
    int cursor = 0, p = pos, newEm = 0;
    bool emit = false, more = on && p < NP;
    double ql = 0, qh = 0;
    if (on) { nGl = 0; nCl = 0; gmask = 0; }
    while (__any_sync(FULL, more)) {
      if (more) {
        const int id = sOrder[p * PE_SLOTS];
        const bool vb = validBit(id);
        const double l = sLo[id * PE_SLOTS], h = sHi[id * PE_SLOTS];
        if (!emit) {
          const bool have = cursor < nEm && sEmPos[cursor * PE_SLOTS] == p;
          const int k = cursor;
          if (have) ++cursor;
          if (vb) {  // valid_ on the whole path: interval = timeRange (body-pair-collision.cc:51-55)
            if (ihi > L) smallest = p;
            ilo = (ilo < 0.0) ? 0.0 : ilo;
            ihi = (L < ihi) ? L : ihi;
            if (have) ++discarded;
          } else if (l <= ilo && ihi <= h) {  // validInterval_ contains the running interval (:56-63)
            if (have) ++discarded;
          } else if (have) {
            const double a = sResAAll[k * PE_SLOTS + lane], b = sResBAll[k * PE_SLOTS + lane];
            if (b == -INF) {  // collision: a carries the geometry pair (computeDistanceLowerBound :186-189)
              finishEdge(false, HPPB200_REASON_COLLISION, t, lastValid, (int)a);
              more = false;
            } else if (!(l > h) && (h < a || b < l)) {
              // validInterval_.insert (:90) would open a second interval: hand the edge to the walker
              ++overflows;
              finishEdge(false, A.overflow_reason, t, lastValid, -1);
              more = false;
            } else {
              double nl = a, nh = b;
              if (!(l > h)) { nl = (a < l) ? a : l; nh = (h > b) ? h : b; }
              sLo[id * PE_SLOTS] = nl;
              sHi[id * PE_SLOTS] = nh;
              if (nl <= 0.0 && L <= nh) sValid[(id >> 5) * PE_SLOTS] |= 1u << (id & 31);  // :92-96
              // validateIntervals: fold into the running intersection
              if (ihi > b) smallest = p;
              ilo = (ilo < a) ? a : ilo;
              ihi = (b < ihi) ? b : ihi;
            }
          } else {
            // the reference evaluates this pair now and its result is not there yet
            emit = true;
            pos = p;
            ql = fmax(ilo, t - hp);
            qh = fmin(ihi, t + hp);
          }
        }
        if (emit && !vb && (p == pos || !(l <= ql && qh <= h))) {
          sEmPos[newEm * PE_SLOTS] = (uint8_t)p;
          if (T.bodypairs[id].has_gjk) { gmask |= 1u << newEm; ++nGl; } else ++nCl;
          ++newEm;
        }
        ++p;
        more = more && p < NP && newEm < PE_RMAX;
      }
    }
    if (on) nEm = newEm;
    if (on && active && !emit) {
      // sample complete.  src/continuous-validation.cc:144-146: smallest interval first
      if (NP > 1 && smallest != 0) {
        const uint8_t a0 = sOrder[0], as = sOrder[smallest * PE_SLOTS];
        sOrder[0] = as;
        sOrder[smallest * PE_SLOTS] = a0;
      }
      // progressive.cc:104-117
      const double tprev = t;
      t = reverse ? ilo : ihi;
      lastValid = tprev;
      hp = A.alpha * fabs(t - tprev);
      if (reverse ? t <= 0.0 : t >= L) { t = reverse ? 0.0 : L; finished++; }
      sampleOpen = false;
      if (finished >= 2) finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
    }
  };

  for (;;) {  // @region mainloop
    if (warp == 0) {
      // =============================== control phase ===============================================
      ++ticks;
      bool on = active && sampleOpen;
      for (;;) {
        scanPass(on);
        // refill: finished lanes pull the next edges (ballot + popc prefix, one atomic per warp)  @region refill
        const bool want = !active && !retired;
        const unsigned m = __ballot_sync(FULL, want);
        if (m) {
          const int leader = __ffs(m) - 1;
          unsigned long long b = 0;
          if (lane == leader) b = atomicAdd(A.counter, (unsigned long long)__popc(m));
          b = __shfl_sync(FULL, b, leader);
          if (want) {
            const unsigned long long idx = b + __popc(m & ((1u << lane) - 1u));
            if (idx >= A.n) retired = true;
            else {
              e = (size_t)idx;
              q1 = A.q1 + e * (size_t)mv.nq;
              q2 = A.q2 + e * (size_t)mv.nq;
              L = A.L[e];
              iters = 0;
              finished = 0;
              lastValid = reverse ? L : 0.0;
              t = lastValid;
              hp = INF;
              if (reverse ? t <= 0.0 : t >= L) finished++;
              for (int id = 0; id < NP; ++id) { sLo[id * PE_SLOTS] = INF; sHi[id * PE_SLOTS] = -INF; sOrder[id * PE_SLOTS] = (uint8_t)id; }
              for (int w = 0; w < SL.validWords; ++w) sValid[w * PE_SLOTS] = 0u;
              sTAll[PE_SLOTS + lane] = L;
              if (L != 0) {
                for (int j = 0; j < mv.njoints; ++j) {
                  const DevJoint& J = T.joints[j];
                  const int nvj = J.type == HPPB200_JOINT_FREEFLYER ? 6 : 1;
                  sS[J.idx_v * PE_SLOTS] = jointVelocityNorm(q1, q2, J.idx_q, nvj, L);
                }
              }
              active = true;
              sampleOpen = false;
            }
          }
        }
        // next sample of every lane that completed one (or just got an edge)
        on = active && !sampleOpen;
        if (!__any_sync(FULL, on)) break;
        if (on) {
          if (iters >= A.max_iter) {
            ++caps;
            finishEdge(false, HPPB200_REASON_ITER_CAP, t, lastValid, -1);
            on = false;
          } else {
            ++iters;
            ++samples;
            pos = 0; smallest = 0; ilo = -INF; ihi = INF; nEm = 0;
            sampleOpen = true;
            fkDone = false;
          }
        }
      }
      // every active lane now waits for the result of the pair at `pos`  @region fk
      if (active && !fkDone) {
        // Path::eval -> StraightPath::impl_compute: stored end points are returned bit-exactly
        if (t == 0.0 || L == 0.0) qmode = 0;
        else if (t == L) qmode = 1;
        else { qmode = 2; u = (t - 0.0) / L; }
        forwardKinematics<TREE>(T, mv, EdgeQ{q1, q2, qmode, u}, scratchCTA + (size_t)lane * slotStride, 1);
        sTAll[lane] = t;
        fkDone = true;
      }
      // one warp prefix sum assigns queue slots: GJK pairs from the front, closed-form pairs from the back  @region emit
      if (!active) { nEm = 0; nGl = 0; nCl = 0; }
      int gOff = nGl, cOff = nCl;  // inclusive scans
      for (int o = 1; o < 32; o <<= 1) {
        const int g2 = __shfl_up_sync(FULL, gOff, o), c2 = __shfl_up_sync(FULL, cOff, o);
        if (lane >= o) { gOff += g2; cOff += c2; }
      }
      const int totG = __shfl_sync(FULL, gOff, 31), totC = __shfl_sync(FULL, cOff, 31);
      gOff -= nGl; cOff -= nCl;
      const int maxEm = __reduce_max_sync(FULL, nEm);
      for (int k = 0; k < maxEm; ++k) {
        if (k < nEm) {
          const int id = sOrder[(int)sEmPos[k * PE_SLOTS] * PE_SLOTS];
          const uint32_t job = (uint32_t)lane | ((uint32_t)k << 5) | ((uint32_t)id << 9);
          if ((gmask >> k) & 1u) queue[gOff++] = job;
          else queue[PE_QCAP - 1 - (cOff++)] = job;
        }
      }
      const bool alive = __any_sync(FULL, active);
      if (lane == 0) { ctl->nG = totG; ctl->nC = totC; ctl->head = 0; ctl->done = alive ? 0 : 1; }
    }
    __syncthreads();
    if (ctl->done) break;
    // ================================= narrow phase ================================================  // @region narrow
    // One body pair per thread: BodyPairCollision::validateConfiguration :65-89 up to the interval
    // [t - (hd + ht), t + (hd + ht)]; a collision is reported as (geometry pair, -inf).
    {
      const int nG = ctl->nG, total = nG + ctl->nC;
      for (;;) {
        int i0 = 0;
        if (lane == 0) i0 = atomicAdd(&ctl->head, 32);
        i0 = __shfl_sync(FULL, i0, 0);
        if (i0 >= total) break;
        const int i = i0 + lane;
        if (i < total) {
          const uint32_t job = i < nG ? queue[i] : queue[PE_QCAP - 1 - (i - nG)];
          const int slot = job & 31, k = (job >> 5) & 15, id = job >> 9;
          const DevBodyPair& B = T.bodypairs[id];
          const double* store = scratchCTA + (size_t)slot * slotStride;
          double dlb = INF;
          int hitGp = -1;
          for (int gp = B.pair_begin; gp < B.pair_end; ++gp) {
            const DevPair pr = T.bgpairs[gp];
            const double d = pairLowerBound(T, pr, store, 1, mv.vertices, cnt);
            ++tests;
            if (d <= HPPB200_COLLISION_THRESHOLD) { hitGp = gp; break; }
            if (d < dlb) dlb = d;
          }
          double ra, rb;
          if (hitGp >= 0) { ra = (double)hitGp; rb = -INF; }
          else {
            const double tt = sTAll[slot], LL = sTAll[PE_SLOTS + slot];
            double hd, ht;  // :72-85
            if (dlb == INF) { hd = INF; ht = 0; }
            else {
              // SolidSolidCollision::computeMaximalVelocity (solid-solid-collision.cc:76-99) on the per-joint
              // norms of Vb = |q2 - q1| / L (StraightPath::impl_velocityBound, straight-path.cc:132-140)
              double Vm;
              if (LL == 0) Vm = INF;  // BodyPairCollision::setupPath, zero-length path
              else {
                const double* S = sSAll + slot;
                double ab = 0;
                for (int c = B.coef_begin; c < B.coef_end; ++c) ab += T.coefs[c].value * S[T.coefs[c].idx_v * PE_SLOTS];
                Vm = ab;
                if (B.joint_b != 0) {
                  double ba = 0;
                  for (int c = B.rcoef_begin; c < B.rcoef_end; ++c) ba += T.rcoefs[c].value * S[T.rcoefs[c].idx_v * PE_SLOTS];
                  if (ba < ab) Vm = ba;
                }
              }
              hd = dlb / Vm;
              ht = (Vm != 0) ? 2 * A.tol / Vm : INF;
            }
            ra = tt - (hd + ht);
            rb = tt + (hd + ht);
          }
          sResAAll[k * PE_SLOTS + slot] = ra;
          sResBAll[k * PE_SLOTS + slot] = rb;
        }
      }
    }
    __syncthreads();
  }

  unsigned ts = __reduce_add_sync(FULL, tests);  // @region stats
  unsigned gi = __reduce_add_sync(FULL, cnt.iterations);
  unsigned gc = __reduce_add_sync(FULL, cnt.cap_hits);
  unsigned sm = __reduce_add_sync(FULL, samples);
  unsigned cp = __reduce_add_sync(FULL, caps);
  unsigned ov = __reduce_add_sync(FULL, overflows);
  unsigned dc = __reduce_add_sync(FULL, discarded);
  if (lane == 0 && stats) {
    if (ts) atomicAdd(&stats->pair_tests, (unsigned long long)ts);
    if (gi) atomicAdd(&stats->gjk_iterations, (unsigned long long)gi);
    if (gc) atomicAdd(&stats->gjk_cap_hits, (unsigned long long)gc);
    if (sm) atomicAdd(&stats->edge_samples, (unsigned long long)sm);
    if (cp) atomicAdd(&stats->iter_cap_hits, (unsigned long long)cp);
    if (ov) atomicAdd(&stats->overflow, (unsigned long long)ov);
    if (dc) atomicAdd(&stats->discarded_tests, (unsigned long long)dc);
    if (ticks) atomicAdd(&stats->engine_ticks, (unsigned long long)ticks);
  }
}

// Returns false when the model does not fit this engine (the caller then uses the generic walker).
template <bool TREE, int THREADS, int MINB>
static bool launchPE(Workspace* ws, const PELayout& SL, PEArgs A, size_t n) {
  const ModelView& mv = ws->view;
  auto kern = progressiveEngineKernel<TREE, THREADS, MINB>;
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SL.total));
  int perSM = 0;
  HPPB200_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kern, THREADS, SL.total));
  if (perSM < 1) return false;
  size_t grid = (size_t)ws->num_sms * perSM;
  const size_t needed = (n + PE_SLOTS - 1) / PE_SLOTS;
  if (grid > needed) grid = needed;
  const size_t scratchDoubles = grid * PE_SLOTS * (size_t)A.fkSlots * 12;
  ws->d_state.reserve(256 + scratchDoubles * sizeof(double));
  A.counter = (unsigned long long*)ws->d_state.p;
  A.scratch = (double*)(ws->d_state.p + 256);
  HPPB200_CUDA_CHECK(cudaMemsetAsync(A.counter, 0, sizeof(unsigned long long), ws->stream));
  kern<<<(unsigned)grid, THREADS, SL.total, ws->stream>>>(mv, A, SL, ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  return true;
}

bool launchProgressiveEngine(Workspace* ws, double tol, int reverse, const double* d_q1, const double* d_q2,
                             const double* d_L, size_t n, uint8_t* d_valid, double* d_tfail, double* d_tlast,
                             int32_t* d_pair, uint8_t* d_reason, int overflow_reason) {
  const ModelView& mv = ws->view;
  const Model* m = ws->model;
  if (mv.nbodypairs > 255 || mv.nbodypairs < 1) return false;
  const PELayout SL = peLayout(mv);
  if (SL.total > 110u * 1024u) return false;  // at least two CTAs per SM
  const bool tree = modelIsTree(m);
  PEArgs A{};
  A.q1 = d_q1; A.q2 = d_q2; A.L = d_L; A.n = n; A.reverse = reverse; A.tol = tol; A.max_iter = ws->max_iterations;
  A.valid = d_valid; A.t_fail = d_tfail; A.t_last = d_tlast; A.pair = d_pair; A.reason = d_reason;
  A.fkSlots = mv.ngeoms + (tree ? mv.njoints : 0);
  A.overflow_reason = overflow_reason;
  A.alpha = ws->engine_alpha;
  if (ws->engine_threads == 64) return tree ? launchPE<true, 64, 4>(ws, SL, A, n) : launchPE<false, 64, 4>(ws, SL, A, n);
  if (ws->engine_threads == 96) return tree ? launchPE<true, 96, 4>(ws, SL, A, n) : launchPE<false, 96, 4>(ws, SL, A, n);
  return tree ? launchPE<true, 128, 3>(ws, SL, A, n) : launchPE<false, 128, 3>(ws, SL, A, n);
}

}  // namespace hppb200
