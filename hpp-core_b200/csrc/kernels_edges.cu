// Batched StraightPath (edge) validation: pathValidation::Discretized
// (reference src/path-validation/discretized.cc:55-107), continuousValidation::Progressive
// (src/continuous-validation/progressive.cc:78-129) and ::Dichotomy
// (src/continuous-validation/dichotomy.cc:68-132), with the per-sample logic of
// ContinuousValidation::validateConfiguration (src/continuous-validation.cc:128-148),
// validateIntervals (include/hpp/core/continuous-validation.hh:267-298) and
// BodyPairCollision::validateConfiguration (src/continuous-validation/body-pair-collision.cc:45-99),
// the velocity bound of SolidSolidCollision::computeMaximalVelocity
// (src/continuous-validation/solid-solid-collision.cc:76-99) and StraightPath::{impl_compute,
// impl_velocityBound} (src/straight-path.cc:91-140).
//
// Execution model ("edge walker"): a persistent grid; every LANE owns one edge at a time and
// walks its samples exactly like the reference's sequential loop, so skip rules and the
// smallest-interval-first reordering are reproduced per edge (canonical behaviour: pair order
// is reset to construction order at the start of every edge).
// Lanes of a warp stay in lock-step on the expensive step only: in every trip of the main loop
// each lane that has a pending geometry pair evaluates exactly ONE narrow-phase test; the cheap
// skip tests (interval containment) are fast-forwarded per lane in between.  When an edge ends
// the lane pulls the next edge from a global counter (warp-aggregated atomic: ballot + popc
// prefix), i.e. finished edges are compacted out and new work is queued without leaving the
// kernel.  Per-lane state that does not fit registers (geometry placements, per-body-pair
// validated interval sets, pair order) lives in global scratch laid out SoA over lanes
// (coalesced, L2-resident).
#include "dev_model.cuh"

namespace hppb200 {

constexpr int EDGE_THREADS = 128;

// scratch layout (element offsets; every field is [index][nslots])
struct EdgeScratch {
  double* store;      // [(ngeoms(+njoints))*12][nslots]
  double* ivlo;       // [nbodypairs*K][nslots]
  double* ivhi;
  uint8_t* ivcnt;     // [nbodypairs][nslots]
  uint8_t* ivvalid;   // [nbodypairs][nslots]
  uint16_t* order;    // [nbodypairs][nslots]
  double* sublo;      // [KS][nslots]  (Dichotomy validSubset)
  double* subhi;
  size_t nslots;
  int K, KS;
};

struct EdgeArgs {
  const double* q1;
  const double* q2;
  const double* L;
  size_t n;
  uint32_t flags;
  int reverse;
  double param;  // step size / tolerance
  uint32_t max_iter;
  uint8_t* valid;
  double* t_fail;
  double* t_last;
  int32_t* pair;
  uint8_t* reason;
  unsigned long long* counter;  // next edge to hand out
  const uint32_t* edge_list;    // optional indirection (second pass over overflowed edges)
  int overflow_reason;          // reason written when an interval list overflows
};

// Sorted disjoint closed-interval list with capacity cap, SoA-strided.  Same observable
// behaviour as the reference Intervals::unionInterval (src/continuous-validation/intervals.hh:53-132)
// and boost::icl::interval_set<closed>::insert used in body-pair-collision.cc:90.
// Returns false on capacity overflow.
__device__ __forceinline__ bool ivUnion(double* lo, double* hi, size_t stride, int& cnt, int cap, double a, double b) {
  int i = 0;
  while (i < cnt && hi[i * stride] < a) ++i;
  if (i == cnt) {
    if (cnt == cap) return false;
    lo[i * stride] = a; hi[i * stride] = b; ++cnt;
    return true;
  }
  if (b < lo[i * stride]) {
    if (cnt == cap) return false;
    for (int k = cnt; k > i; --k) { lo[k * stride] = lo[(k - 1) * stride]; hi[k * stride] = hi[(k - 1) * stride]; }
    lo[i * stride] = a; hi[i * stride] = b; ++cnt;
    return true;
  }
  double nlo = lo[i * stride];
  if (a < nlo) nlo = a;
  double nhi = b;
  int j = i;
  while (j < cnt && lo[j * stride] <= b) {
    const double h = hi[j * stride];
    if (h > nhi) nhi = h;
    ++j;
  }
  lo[i * stride] = nlo; hi[i * stride] = nhi;
  const int removed = j - i - 1;
  if (removed > 0) {
    for (int k = i + 1; k + removed < cnt; ++k) { lo[k * stride] = lo[(k + removed) * stride]; hi[k * stride] = hi[(k + removed) * stride]; }
    cnt -= removed;
  }
  return true;
}
__device__ __forceinline__ bool ivContains(const double* lo, const double* hi, size_t stride, int cnt, double a, double b) {
  for (int i = 0; i < cnt; ++i)
    if (lo[i * stride] <= a && b <= hi[i * stride]) return true;
  return false;
}

enum { PH_NEED_EDGE = 0, PH_SAMPLE = 1, PH_PAIRS = 2, PH_RETIRED = 3 };

template <int METHOD, bool TREE>
__global__ void __launch_bounds__(EDGE_THREADS)
edgeWalkerKernel(ModelView mv, EdgeArgs A, EdgeScratch S, DeviceStats* stats) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint64_t* bar = (uint64_t*)smem;
  unsigned char* sblob = smem + 16;
  stageBlobTMA(sblob, mv.blob, mv.blob_bytes, bar);
  const Tables T = makeTables(sblob, mv);

  constexpr bool DISC = METHOD == HPPB200_METHOD_DISCRETIZED;
  constexpr bool PROG = METHOD == HPPB200_METHOD_PROGRESSIVE;
  constexpr bool DICH = METHOD == HPPB200_METHOD_DICHOTOMY;
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const size_t NS = S.nslots;
  const size_t slot = (size_t)blockIdx.x * EDGE_THREADS + threadIdx.x;
  double* store = S.store + slot;
  const bool reverse = A.reverse != 0;
  const int NP = DISC ? mv.ncpairs : mv.nbodypairs;
  const double INF = CUDART_INF;

  // ---- lane state -----------------------------------------------------------------
  int phase = PH_NEED_EDGE;
  size_t e = 0;
  const double* q1 = nullptr;
  const double* q2 = nullptr;
  double L = 0, t = 0, lastValid = 0, u = 0;
  int qmode = 0;  // 0: initial config, 1: end config, 2: interpolate
  int finished = 0;
  unsigned iters = 0;
  int pos = 0, gp = -1, gpEnd = 0, curId = 0, smallest = 0, subCnt = 0;
  double ilo = 0, ihi = 0, dlb = 0;
  bool pending = false;
  GjkCounters cnt = {0u, 0u, 0u};
  unsigned tests = 0, samples = 0, caps = 0, overflows = 0;

  auto qAt = [&](int k) -> double {
    const double a = q1[k];
    if (qmode == 0) return a;
    const double b = q2[k];
    if (qmode == 1) return b;
    return fma(u, b - a, a);  // LiegroupSpace::interpolate on R^n
  };
  auto finishEdge = [&](bool ok, int rs, double tf, double tl, int pr) {
    A.valid[e] = ok ? 1 : 0;
    if (A.t_fail) A.t_fail[e] = tf;
    if (A.t_last) A.t_last[e] = tl;
    if (A.pair) A.pair[e] = pr;
    if (A.reason) A.reason[e] = (uint8_t)rs;
    phase = PH_NEED_EDGE;
    pending = false;
  };
  // SolidSolidCollision::computeMaximalVelocity on Vb = |q2 - q1| / L (StraightPath::impl_velocityBound)
  auto maximalVelocity = [&](const DevBodyPair& B) -> double {
    if (L == 0) return INF;  // BodyPairCollision::setupPath, zero-length path
    auto sum = [&](const DevCoef* c, int b, int en) {
      double acc = 0;
      for (int i = b; i < en; ++i) {
        double s = 0;
        for (int k = 0; k < c[i].nv; ++k) {
          const double vb = fabs(q2[c[i].idx_v + k] - q1[c[i].idx_v + k]) / L;
          s = fma(vb, vb, s);
        }
        acc += c[i].value * sqrt(s);
      }
      return acc;
    };
    const double ab = sum(T.coefs, B.coef_begin, B.coef_end);
    if (B.joint_b != 0) {
      const double ba = sum(T.rcoefs, B.rcoef_begin, B.rcoef_end);
      if (ba < ab) return ba;
    }
    return ab;
  };
  // validateIntervals: fold one body pair's interval into the running intersection
  auto applyInterval = [&](double tlo, double thi) {
    if (ihi > thi) smallest = pos;
    ilo = (ilo < tlo) ? tlo : ilo;
    ihi = (thi < ihi) ? thi : ihi;
  };

  for (;;) {
    // ================= advance the lane until it has one pending narrow-phase test ========
    for (int round = 0; round < 3 && !pending && phase != PH_RETIRED; ++round) {
      if (phase == PH_NEED_EDGE) {
        // warp-aggregated fetch of the next edges (ballot + popc prefix)
        const unsigned m = __activemask();
        const int leader = __ffs(m) - 1;
        unsigned long long b = 0;
        if (lane == leader) b = atomicAdd(A.counter, (unsigned long long)__popc(m));
        b = __shfl_sync(m, b, leader);
        const unsigned long long idx = b + __popc(m & ((1u << lane) - 1u));
        if (idx >= A.n) { phase = PH_RETIRED; break; }
        e = A.edge_list ? (size_t)A.edge_list[idx] : (size_t)idx;
        q1 = A.q1 + e * (size_t)mv.nq;
        q2 = A.q2 + e * (size_t)mv.nq;
        L = A.L[e];
        iters = 0;
        finished = 0;
        lastValid = reverse ? L : 0.0;
        t = lastValid;
        if (PROG) { if (reverse ? t <= 0.0 : t >= L) finished++; }
        if (!DISC) {
          for (int id = 0; id < NP; ++id) {
            S.ivcnt[id * NS + slot] = 0;
            S.ivvalid[id * NS + slot] = 0;
            S.order[id * NS + slot] = (uint16_t)id;
          }
        }
        if (DICH) { S.sublo[slot] = t; S.subhi[slot] = t; subCnt = 1; }
        phase = PH_SAMPLE;
      }
      if (phase == PH_SAMPLE) {
        if (iters >= A.max_iter) {
          ++caps;
          double tl = lastValid;
          if (DICH) tl = reverse ? S.sublo[(size_t)(subCnt - 1) * NS + slot] : S.subhi[slot];
          finishEdge(false, HPPB200_REASON_ITER_CAP, t, tl, -1);
          continue;
        }
        ++iters;
        ++samples;
        // Path::eval -> StraightPath::impl_compute: stored end points are returned bit-exactly
        if (t == 0.0 || L == 0.0) qmode = 0;
        else if (t == L) qmode = 1;
        else { qmode = 2; u = (t - 0.0) / L; }
        if (DISC && (A.flags & HPPB200_CHECK_JOINT_BOUNDS)) {
          int bad = -1;
          for (int k = 0; k < mv.nq && bad < 0; ++k) {
            const double lo = T.lower[k], hi = T.upper[k];
            if (lo > hi) continue;
            const double v = qAt(k);
            if (v < lo || hi < v) bad = k;
          }
          if (bad >= 0) { finishEdge(false, HPPB200_REASON_JOINT_BOUND, t, lastValid, bad); continue; }
        }
        forwardKinematics<TREE>(T, mv, qAt, store, NS);
        pos = 0;
        gp = -1;
        smallest = 0;
        ilo = -INF;
        ihi = INF;
        phase = PH_PAIRS;
      }
      if (phase == PH_PAIRS) {
        if (DISC) {
          if (pos < NP) { pending = true; break; }
        } else {
          // fast-forward over body pairs the reference would skip (body-pair-collision.cc:51-63)
          while (pos < NP) {
            const int id = S.order[pos * NS + slot];
            if (S.ivvalid[id * NS + slot]) { applyInterval(0.0, L); ++pos; continue; }
            const size_t o = ((size_t)id * S.K) * NS + slot;
            if (ivContains(S.ivlo + o, S.ivhi + o, NS, S.ivcnt[id * NS + slot], ilo, ihi)) { ++pos; continue; }
            curId = id;
            gp = T.bodypairs[id].pair_begin;
            gpEnd = T.bodypairs[id].pair_end;
            dlb = INF;
            break;
          }
          if (pos < NP) { pending = true; break; }
        }
        // ---- end of a collision-free sample ----
        if (DISC) {
          lastValid = t;
          t += reverse ? -A.param : A.param;
          if ((reverse && t < 0.0) || (!reverse && t > L)) { t = reverse ? 0.0 : L; finished++; }
          if (finished < 2) phase = PH_SAMPLE;
          else finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
        } else {
          // src/continuous-validation.cc:144-146: smallest interval first
          if (NP > 1 && smallest != 0) {
            const uint16_t a0 = S.order[slot], as = S.order[smallest * NS + slot];
            S.order[slot] = as;
            S.order[smallest * NS + slot] = a0;
          }
          if (PROG) {
            const double tprev = t;
            t = reverse ? ilo : ihi;
            lastValid = tprev;
            if (reverse ? t <= 0.0 : t >= L) { t = reverse ? 0.0 : L; finished++; }
            if (finished < 2) phase = PH_SAMPLE;
            else finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
          } else {
            if (!ivUnion(S.sublo + slot, S.subhi + slot, NS, subCnt, S.KS, ilo, ihi)) {
              if (A.overflow_reason == HPPB200_REASON_ITER_CAP) ++caps; else ++overflows;
              finishEdge(false, A.overflow_reason, t, reverse ? S.sublo[(size_t)(subCnt - 1) * NS + slot] : S.subhi[slot], -1);
              continue;
            }
            if (ivContains(S.sublo + slot, S.subhi + slot, NS, subCnt, 0.0, L)) {
              finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
            } else {
              // middle of the first non-validated gap (dichotomy.cc:103-109)
              double t0, t1;
              if (!reverse) {
                t0 = S.subhi[slot];
                t1 = subCnt > 1 ? S.sublo[NS + slot] : L;
              } else {
                t0 = S.sublo[(size_t)(subCnt - 1) * NS + slot];
                t1 = subCnt > 1 ? S.subhi[(size_t)(subCnt - 2) * NS + slot] : 0.0;
              }
              t = .5 * (t0 + t1);
              phase = PH_SAMPLE;
            }
          }
        }
      }
    }
    if (__all_sync(FULL, phase == PH_RETIRED)) break;

    // ================= one narrow-phase test per lane, warp in lock-step ==================
    if (pending) {
      const DevPair pr = DISC ? T.cpairs[pos] : T.bgpairs[gp];
      const double d = pairLowerBound(T, pr, store, NS, mv.vertices, cnt);
      ++tests;
      if (d <= HPPB200_COLLISION_THRESHOLD) {
        double tl = lastValid;
        if (DICH) tl = reverse ? S.sublo[(size_t)(subCnt - 1) * NS + slot] : S.subhi[slot];
        finishEdge(false, HPPB200_REASON_COLLISION, t, tl, DISC ? pos : gp);
      } else if (DISC) {
        ++pos;
        pending = false;
      } else {
        if (d < dlb) dlb = d;
        ++gp;
        if (gp >= gpEnd) {
          // BodyPairCollision::validateConfiguration :72-96
          const DevBodyPair& B = T.bodypairs[curId];
          double hd, ht;
          if (dlb == INF) { hd = INF; ht = 0; }
          else {
            const double Vm = maximalVelocity(B);
            hd = dlb / Vm;
            ht = (Vm != 0) ? 2 * A.param / Vm : INF;
          }
          const double a = t - (hd + ht), b = t + (hd + ht);
          const size_t o = ((size_t)curId * S.K) * NS + slot;
          int c = S.ivcnt[curId * NS + slot];
          if (!ivUnion(S.ivlo + o, S.ivhi + o, NS, c, S.K, a, b)) {
            if (A.overflow_reason == HPPB200_REASON_ITER_CAP) ++caps; else ++overflows;
            double tl = lastValid;
            if (DICH) tl = reverse ? S.sublo[(size_t)(subCnt - 1) * NS + slot] : S.subhi[slot];
            finishEdge(false, A.overflow_reason, t, tl, -1);
          } else {
            S.ivcnt[curId * NS + slot] = (uint8_t)c;
            if (ivContains(S.ivlo + o, S.ivhi + o, NS, c, 0.0, L)) S.ivvalid[curId * NS + slot] = 1;
            applyInterval(a, b);
            ++pos;
            gp = -1;
            pending = false;
          }
        }
        // else: same body pair, next geometry pair stays pending
      }
    }
  }

  unsigned ts = __reduce_add_sync(FULL, tests);
  unsigned gi = __reduce_add_sync(FULL, cnt.iterations);
  unsigned gc = __reduce_add_sync(FULL, cnt.cap_hits);
  unsigned sm = __reduce_add_sync(FULL, samples);
  unsigned cp = __reduce_add_sync(FULL, caps);
  unsigned ov = __reduce_add_sync(FULL, overflows);
  if (lane == 0 && stats) {
    if (ts) atomicAdd(&stats->pair_tests, (unsigned long long)ts);
    if (gi) atomicAdd(&stats->gjk_iterations, (unsigned long long)gi);
    if (gc) atomicAdd(&stats->gjk_cap_hits, (unsigned long long)gc);
    if (sm) atomicAdd(&stats->edge_samples, (unsigned long long)sm);
    if (cp) atomicAdd(&stats->iter_cap_hits, (unsigned long long)cp);
    if (ov) atomicAdd(&stats->overflow, (unsigned long long)ov);
  }
}

// L = WeighedDistance(q1, q2) (reference src/weighed-distance.cc:272-289,
// steeringMethod::Straight::impl_compute src/steering-method/straight.cc:38-51)
__global__ void straightLengthsKernel(ModelView mv, const double* __restrict__ q1, const double* __restrict__ q2, size_t n,
                                      double* __restrict__ L) {
  const Tables T = makeTables(mv.blob, mv);
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* a = q1 + i * (size_t)mv.nq;
  const double* b = q2 + i * (size_t)mv.nq;
  double res = 0;
  for (int j = 0; j < mv.njoints; ++j) {
    const DevJoint& J = T.joints[j];
    const double w = T.weights[j];
    const double w2 = w * w;
    const double dq = b[J.idx_q] - a[J.idx_q];
    double d;
    if (J.type <= HPPB200_JOINT_REVOLUTE_Z) d = w2 * (dq * dq);
    else d = dq * dq;
    res += d;
  }
  L[i] = sqrt(res);
}

void launchStraightLengths(Workspace* ws, const double* d_q1, const double* d_q2, size_t n, double* d_L) {
  if (n == 0) return;
  const unsigned grid = (unsigned)((n + 255) / 256);
  straightLengthsKernel<<<grid, 256, 0, ws->stream>>>(ws->view, d_q1, d_q2, n, d_L);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
}

constexpr int REASON_OVERFLOW_INTERNAL = 4;  // never leaves the library

__global__ void collectOverflowKernel(const uint8_t* __restrict__ reason, size_t n, uint32_t* __restrict__ list,
                                      unsigned long long* __restrict__ count) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && reason[i] == REASON_OVERFLOW_INTERNAL) list[atomicAdd(count, 1ull)] = (uint32_t)i;
}

template <int METHOD>
static void launchEdgeWalker(Workspace* ws, const EdgeArgs& A0, int K, int KS, size_t maxSlots) {
  const bool tree = modelIsTree(ws->model);
  auto kern = tree ? edgeWalkerKernel<METHOD, true> : edgeWalkerKernel<METHOD, false>;
  const ModelView& mv = ws->view;
  const size_t smemBytes = 16 + mv.blob_bytes;
  HPPB200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
  int perSM = 0;
  HPPB200_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kern, EDGE_THREADS, smemBytes));
  if (perSM < 1) throw std::runtime_error("edge kernel does not fit on an SM");
  size_t grid = (size_t)ws->num_sms * perSM;
  const size_t needed = (A0.n + EDGE_THREADS - 1) / EDGE_THREADS;
  if (grid > needed) grid = needed;
  if (grid * EDGE_THREADS > maxSlots) grid = std::max<size_t>(1, maxSlots / EDGE_THREADS);
  const size_t NS = grid * EDGE_THREADS;

  // carve the scratch (8-byte aligned fields)
  const size_t fkSlots = (size_t)mv.ngeoms + (tree ? mv.njoints : 0);
  const size_t nbp = METHOD == HPPB200_METHOD_DISCRETIZED ? 0 : (size_t)mv.nbodypairs;
  auto al = [](size_t x) { return (x + 255) & ~size_t(255); };
  size_t off = 0;
  const size_t o_counter = off; off += 256;
  const size_t o_store = off; off += al(fkSlots * 12 * NS * sizeof(double));
  const size_t o_ivlo = off; off += al(nbp * K * NS * sizeof(double));
  const size_t o_ivhi = off; off += al(nbp * K * NS * sizeof(double));
  const size_t o_sublo = off; off += al((size_t)KS * NS * sizeof(double));
  const size_t o_subhi = off; off += al((size_t)KS * NS * sizeof(double));
  const size_t o_order = off; off += al(nbp * NS * sizeof(uint16_t));
  const size_t o_cnt = off; off += al(nbp * NS);
  const size_t o_valid = off; off += al(nbp * NS);
  ws->d_state.reserve(off);
  unsigned char* base = ws->d_state.p;
  EdgeScratch S;
  S.store = (double*)(base + o_store);
  S.ivlo = (double*)(base + o_ivlo);
  S.ivhi = (double*)(base + o_ivhi);
  S.sublo = (double*)(base + o_sublo);
  S.subhi = (double*)(base + o_subhi);
  S.order = (uint16_t*)(base + o_order);
  S.ivcnt = base + o_cnt;
  S.ivvalid = base + o_valid;
  S.nslots = NS;
  S.K = K;
  S.KS = KS;
  EdgeArgs A = A0;
  A.counter = (unsigned long long*)(base + o_counter);

  HPPB200_CUDA_CHECK(cudaMemsetAsync(A.counter, 0, sizeof(unsigned long long), ws->stream));
  kern<<<(unsigned)grid, EDGE_THREADS, smemBytes, ws->stream>>>(mv, A, S, ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
}

void launchValidateEdges(Workspace* ws, int method, double param, int reverse, const double* d_q1, const double* d_q2,
                         const double* d_L, size_t n, uint32_t flags, uint8_t* d_valid, double* d_tfail, double* d_tlast,
                         int32_t* d_pair, uint8_t* d_reason) {
  if (n == 0) return;
  if (n > 0xffffffffull) throw std::invalid_argument("at most 2^32-1 edges per call");
  if (ws->view.nbodypairs > 65535) throw std::runtime_error("too many body pairs");
  if (!d_reason) {  // the overflow protocol needs a reason array
    ws->d_reason_int.reserve(n);
    d_reason = ws->d_reason_int.p;
  }
  EdgeArgs A{};
  A.q1 = d_q1; A.q2 = d_q2; A.L = d_L; A.n = n; A.flags = flags; A.reverse = reverse; A.param = param;
  A.max_iter = ws->max_iterations;
  A.valid = d_valid; A.t_fail = d_tfail; A.t_last = d_tlast; A.pair = d_pair; A.reason = d_reason;
  A.edge_list = nullptr;
  A.overflow_reason = REASON_OVERFLOW_INTERNAL;
  // Interval-set capacities.  Progressive provably keeps ONE connected validated interval per
  // body pair; Dichotomy keeps short lists.  An edge whose lists overflow the
  // fast-path capacity is re-run ON THE DEVICE with large lists (second pass); if even that
  // overflows it is reported as HPPB200_REASON_ITER_CAP, never silently accepted.
  DeviceStats before, after;
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&before, ws->d_stats, sizeof(before), cudaMemcpyDeviceToHost, ws->stream));
  const size_t allSlots = ~size_t(0);
  if (method == HPPB200_METHOD_DISCRETIZED) { launchEdgeWalker<HPPB200_METHOD_DISCRETIZED>(ws, A, 1, 1, allSlots); return; }
  if (method == HPPB200_METHOD_PROGRESSIVE) launchEdgeWalker<HPPB200_METHOD_PROGRESSIVE>(ws, A, 2, 1, allSlots);
  else launchEdgeWalker<HPPB200_METHOD_DICHOTOMY>(ws, A, 12, 96, allSlots);
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&after, ws->d_stats, sizeof(after), cudaMemcpyDeviceToHost, ws->stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  const unsigned long long nover = after.overflow - before.overflow;
  if (nover == 0) return;
  // second pass: gather the overflowed edges and walk them with large interval lists
  ws->d_list.reserve(nover + 1);
  ws->d_count.reserve(1);
  HPPB200_CUDA_CHECK(cudaMemsetAsync(ws->d_count.p, 0, sizeof(unsigned long long), ws->stream));
  collectOverflowKernel<<<(unsigned)((n + 255) / 256), 256, 0, ws->stream>>>(d_reason, n, ws->d_list.p, ws->d_count.p);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  A.edge_list = ws->d_list.p;
  A.n = (size_t)nover;
  A.overflow_reason = HPPB200_REASON_ITER_CAP;
  const size_t slots2 = 2048;  // 255-entry lists: ~4 KB per body pair per slot
  if (method == HPPB200_METHOD_PROGRESSIVE) launchEdgeWalker<HPPB200_METHOD_PROGRESSIVE>(ws, A, 255, 1, slots2);
  else launchEdgeWalker<HPPB200_METHOD_DICHOTOMY>(ws, A, 255, 2048, slots2);
}

}  // namespace hppb200
