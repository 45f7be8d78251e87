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
// The warp is phase-synchronous: (A) lanes whose edge ended pull the next edges from a global
// counter (warp-aggregated atomic: ballot + popc prefix), i.e. finished edges are compacted out
// and new work is queued without leaving the kernel; (B) every lane sets up its next sample and
// runs FK, all in step; (C) rounds in which each lane fast-forwards over the pairs the reference
// would skip (cheap interval containment tests) and then every lane that has a pending geometry
// pair evaluates exactly ONE narrow-phase test, in step; (D) per-lane end-of-sample update.
// Discretized samples reuse the configuration kernel's exact culling (closed forms + conservative
// bounds decided in a uniform scan, surviving GJK pairs drained in lock-step rounds).
// Per-lane interval sets and pair order live in shared memory when they fit (Progressive:
// one interval per body pair), else in global scratch laid out SoA over lanes; geometry
// placements always live in that scratch (coalesced, L2-resident).
#include <cstdio>
#include <cstdlib>

#include "dev_model.cuh"
#include "dev_spline.cuh"

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
  int smem_state;     // per-lane interval sets live in shared memory (fast path) instead of scratch
};

struct EdgeArgs {
  const double* q1;
  const double* q2;
  const double* P;  // spline edges: parameters [n][4][nv] (q1 = base configurations, q2 unused)
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
  // Two-pass execution of the continuous methods:
  //   mode 1: every lane runs ONLY the first sample of its edge (all body pairs are evaluated there,
  //           so the warp is uniform), saves the per-edge state and appends the edge to cont_list;
  //   mode 2: the walker resumes the saved edges from their second sample;  mode 0: walk from scratch.
  int mode;
  double* sv_dbl;               // [2 + 2*NP + 2][n]: t, lastValid, lo[NP], hi[NP], sublo0, subhi0
  uint16_t* sv_order;           // [NP][n]
  uint8_t* sv_flags;            // [1 + 2*NP][n]: finished, cnt[NP], valid[NP]
  uint32_t* cont_list;
  unsigned long long* cont_count;
  size_t n_total;               // stride of the sv_* arrays (edges in the call)
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

// SPLINE: the edges are path::Spline<BernsteinBasis, 3> (dev_spline.cuh) instead of StraightPaths: A.q1 holds the base
// configurations, A.P the parameters; BodyPairCollision refines its velocity bound on sub-intervals (refine_ = true,
// body-pair-collision.cc:119-173).  Everything else -- sampling loops, skip rules, interval sets -- is path-independent.
template <int METHOD, bool TREE, bool SPLINE = false>
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

  // ---- per-lane interval-set state: shared memory when it fits, else global scratch ----
  double* ivlo; double* ivhi; uint8_t* ivcnt; uint8_t* ivvalid; uint16_t* order;
  size_t SS;  // stride between consecutive entries of one lane
  if (S.smem_state) {
    unsigned char* st = sblob + mv.blob_bytes;
    const size_t nb = (size_t)mv.nbodypairs;
    ivlo = (double*)st + threadIdx.x;
    ivhi = (double*)st + nb * S.K * EDGE_THREADS + threadIdx.x;
    order = (uint16_t*)(st + 2 * nb * S.K * EDGE_THREADS * sizeof(double)) + threadIdx.x;
    ivcnt = (uint8_t*)(order - threadIdx.x + nb * EDGE_THREADS) + threadIdx.x;
    ivvalid = ivcnt + nb * EDGE_THREADS;
    SS = EDGE_THREADS;
  } else {
    ivlo = S.ivlo + slot; ivhi = S.ivhi + slot; ivcnt = S.ivcnt + slot; ivvalid = S.ivvalid + slot; order = S.order + slot;
    SS = NS;
  }
  double* sublo = S.sublo + slot;
  double* subhi = S.subhi + slot;

  // ---- lane state -----------------------------------------------------------------
  bool active = false, retired = false;
  size_t e = 0;
  const double* q1 = nullptr;
  const double* q2 = nullptr;
  const double* Pp = nullptr;  // SPLINE: this edge's parameters
  SplineQ sq;                  // SPLINE: the current sample
  double L = 0, t = 0, lastValid = 0, u = 0;
  int qmode = 0;  // 0: initial config, 1: end config, 2: interpolate
  int finished = 0;
  unsigned iters = 0;
  int subCnt = 0;
  GjkCounters cnt = {0u, 0u, 0u};
  unsigned tests = 0, samples = 0, caps = 0, overflows = 0;

  auto finishEdge = [&](bool ok, int rs, double tf, double tl, int pr) {
    A.valid[e] = ok ? 1 : 0;
    if (A.t_fail) A.t_fail[e] = tf;
    if (A.t_last) A.t_last[e] = tl;
    if (A.pair) A.pair[e] = pr;
    if (A.reason) A.reason[e] = (uint8_t)rs;
    active = false;
  };
  auto dichLast = [&]() -> double { return reverse ? sublo[(size_t)(subCnt - 1) * NS] : subhi[0]; };
  // SolidSolidCollision::computeMaximalVelocity on Vb = |q2 - q1| / L (StraightPath::impl_velocityBound)
  auto maximalVelocity = [&](const DevBodyPair& B) -> double {
    if (L == 0) return INF;  // BodyPairCollision::setupPath, zero-length path
    auto sum = [&](const DevCoef* c, int b, int en) {
      double acc = 0;
      for (int i = b; i < en; ++i) {
        acc += c[i].value * jointVelocityNorm(q1, q2, c[i].idx_q, c[i].nv, L);
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

  for (;;) {
    // ================= A. refill: finished lanes pull the next edges (ballot + popc prefix) ==========
    {
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
            e = A.edge_list ? (size_t)A.edge_list[idx] : (size_t)idx;
            q1 = A.q1 + e * (size_t)mv.nq;
            q2 = SPLINE ? nullptr : A.q2 + e * (size_t)mv.nq;
            if (SPLINE) Pp = A.P + e * 4 * (size_t)mv.nq;
            L = A.L[e];
            if (!DISC && A.mode == 2) {
              // resume after the first sample
              const size_t NT = A.n_total;
              iters = 1;
              t = A.sv_dbl[e];
              lastValid = A.sv_dbl[NT + e];
              finished = A.sv_flags[e];
              for (int id = 0; id < NP; ++id) {
                order[id * SS] = A.sv_order[(size_t)id * NT + e];
                ivcnt[id * SS] = A.sv_flags[(size_t)(1 + id) * NT + e];
                ivvalid[id * SS] = A.sv_flags[(size_t)(1 + NP + id) * NT + e];
                const size_t o = ((size_t)id * S.K) * SS;
                ivlo[o] = A.sv_dbl[(size_t)(2 + id) * NT + e];
                ivhi[o] = A.sv_dbl[(size_t)(2 + NP + id) * NT + e];
              }
              if (DICH) { sublo[0] = A.sv_dbl[(size_t)(2 + 2 * NP) * NT + e]; subhi[0] = A.sv_dbl[(size_t)(3 + 2 * NP) * NT + e]; subCnt = 1; }
            } else {
              iters = 0;
              finished = 0;
              lastValid = reverse ? L : 0.0;
              t = lastValid;
              if (PROG) { if (reverse ? t <= 0.0 : t >= L) finished++; }
              if (!DISC) {
                for (int id = 0; id < NP; ++id) { ivcnt[id * SS] = 0; ivvalid[id * SS] = 0; order[id * SS] = (uint16_t)id; }
              }
              if (DICH) { sublo[0] = t; subhi[0] = t; subCnt = 1; }
            }
            active = true;
          }
        }
      }
      if (!__any_sync(FULL, active)) break;
    }

    // ================= B. sample set-up: q(t), joint bounds, FK — all lanes in step ================
    bool work = false;  // this lane has a sample whose pairs must be processed
    if (active) {
      if (iters >= A.max_iter) {
        ++caps;
        finishEdge(false, HPPB200_REASON_ITER_CAP, t, DICH ? dichLast() : lastValid, -1);
      } else {
        ++iters;
        ++samples;
        if (SPLINE) {
          sq = splineQ(q1, Pp, mv.nq, t, L);  // Spline::impl_compute: u clamped to [0, 1], no end-point shortcut
        } else {
          // Path::eval -> StraightPath::impl_compute: stored end points are returned bit-exactly
          if (t == 0.0 || L == 0.0) qmode = 0;
          else if (t == L) qmode = 1;
          else { qmode = 2; u = (t - 0.0) / L; }
        }
        int bad = -1;
        if (DISC && (A.flags & HPPB200_CHECK_JOINT_BOUNDS)) {
          for (int k = 0; k < mv.nq && bad < 0; ++k) {
            const double lo = T.lower[k], hi = T.upper[k];
            if (lo > hi) continue;
            const double v = SPLINE ? sq(k) : EdgeQ{q1, q2, qmode, u}(k);
            if (v < lo || hi < v) bad = k;
          }
        }
        if (bad >= 0) finishEdge(false, HPPB200_REASON_JOINT_BOUND, t, lastValid, bad);
        else work = true;
      }
    }
    if (work) {
      if (SPLINE) forwardKinematics<TREE>(T, mv, sq, store, NS);
      else forwardKinematics<TREE>(T, mv, EdgeQ{q1, q2, qmode, u}, store, NS);
    }

    // ================= C. pairs ======================================================================
    if (DISC) {
      // C1: uniform scan in reference order.  Closed forms are decided exactly; GJK kinds are culled
      // with the conservative bounds of the configuration kernel; survivors wait in a 4-entry list.
      // C2 (drain): one GJK per lane per round, all lanes in step.
      int hit = INT_MAX;
      int pd0 = -1, pd1 = -1, pd2 = -1, pd3 = -1, npend = 0;
      auto drain = [&]() {
        for (;;) {
          const bool has = npend > 0;
          if (!__any_sync(FULL, has)) break;
          if (has) {
            const int p = pd0;
            pd0 = pd1; pd1 = pd2; pd2 = pd3; pd3 = -1; --npend;
            if (p < hit) {
              const DevPair pr = T.cpairs[p];
              const double d = pairLowerBound(T, pr, store, NS, mv.vertices, cnt);
              ++tests;
              if (d <= HPPB200_COLLISION_THRESHOLD) { hit = p; npend = 0; }
            }
          }
        }
      };
      for (int p = 0; p < NP; ++p) {
        if (work && p < hit) {
          const DevPair pr = T.cpairs[p];
          const DevShape& sa = *(const DevShape*)(pr.a >= 0 ? &T.geoms[pr.a].shape : &T.obstacles[-(pr.a + 1)].shape);
          const DevShape& sb = *(const DevShape*)(pr.b >= 0 ? &T.geoms[pr.b].shape : &T.obstacles[-(pr.b + 1)].shape);
          V3 ca, ua, cb, ub;
          if (pr.a >= 0) {
            const double* M = store + (size_t)pr.a * 12 * NS;
            ua = mk(M[2 * NS], M[5 * NS], M[8 * NS]);
            ca = mk(M[9 * NS], M[10 * NS], M[11 * NS]);
            if (sa.type >= SH_BOX) { const DevBound& B = T.geoms[pr.a].bound; ca = rmul(load12s(M, NS), mk(B.bc[0], B.bc[1], B.bc[2])) + ca; }
          } else { const DevObstacle& O = T.obstacles[-(pr.a + 1)]; ca = mk(O.wc[0], O.wc[1], O.wc[2]); ua = mk(O.P[2], O.P[5], O.P[8]); }
          if (pr.b >= 0) {
            const double* M = store + (size_t)pr.b * 12 * NS;
            ub = mk(M[2 * NS], M[5 * NS], M[8 * NS]);
            cb = mk(M[9 * NS], M[10 * NS], M[11 * NS]);
            if (sb.type >= SH_BOX) { const DevBound& B = T.geoms[pr.b].bound; cb = rmul(load12s(M, NS), mk(B.bc[0], B.bc[1], B.bc[2])) + cb; }
          } else { const DevObstacle& O = T.obstacles[-(pr.b + 1)]; cb = mk(O.wc[0], O.wc[1], O.wc[2]); ub = mk(O.P[2], O.P[5], O.P[8]); }
          if (pr.kind <= PK_CAPSULE_CAPSULE) {
            double d;
            if (pr.kind == PK_SPHERE_SPHERE) { V3 dd = ca - cb; d = sqrt(dot(dd, dd)) - sa.p[0] - sb.p[0]; }
            else if (pr.kind == PK_SPHERE_CAPSULE) d = sqrt(pointSegmentDist2(ca, cb, ub, sb.p[1])) - sa.p[0] - sb.p[0];
            else {
              V3 p1 = axpy(-sa.p[1], ua, ca), q1s = axpy(sa.p[1], ua, ca);
              V3 p2 = axpy(-sb.p[1], ub, cb), q2s = axpy(sb.p[1], ub, cb);
              d = sqrt(segmentSegmentDist2(p1, q1s, p2, q2s)) - sa.p[0] - sb.p[0];
            }
            ++tests;
            if (d - pr.margin <= HPPB200_COLLISION_THRESHOLD) hit = p;
          } else {
            double d2;
            if (pr.bound == 1) d2 = pointSegmentDist2(cb, ca, ua, sa.p[1]);
            else { V3 dd = ca - cb; d2 = dot(dd, dd); }
            bool survive = !(d2 > pr.cull2);
            if (survive && pr.b < 0 && sb.type == SH_BOX) {
              const DevObstacle& O = T.obstacles[-(pr.b + 1)];
              const T3 TB = load12(O.P);
              const double h = pr.bound == 1 ? sa.p[1] : 0.0;
              const V3 e0 = rtmul(TB, axpy(-h, ua, ca) - TB.t), e1 = rtmul(TB, axpy(h, ua, ca) - TB.t);
              const double sx = fmax(fmin(e0.x, e1.x) - sb.p[0], -fmax(e0.x, e1.x) - sb.p[0]);
              const double sy = fmax(fmin(e0.y, e1.y) - sb.p[1], -fmax(e0.y, e1.y) - sb.p[1]);
              const double sz = fmax(fmin(e0.z, e1.z) - sb.p[2], -fmax(e0.z, e1.z) - sb.p[2]);
              const double ra = pr.a >= 0 ? T.geoms[pr.a].bound.br : T.obstacles[-(pr.a + 1)].bound.br;
              if (fmax(sx, fmax(sy, sz)) - ra - pr.margin > 1e-9) survive = false;
            }
            if (survive) {
              if (npend == 0) pd0 = p; else if (npend == 1) pd1 = p; else if (npend == 2) pd2 = p; else pd3 = p;
              ++npend;
            }
          }
        }
        if (__any_sync(FULL, npend == 4)) drain();
      }
      drain();
      if (work) {
        if (hit != INT_MAX) finishEdge(false, HPPB200_REASON_COLLISION, t, lastValid, hit);
        else {
          lastValid = t;
          t += reverse ? -A.param : A.param;
          if ((reverse && t < 0.0) || (!reverse && t > L)) { t = reverse ? 0.0 : L; finished++; }
          if (finished >= 2) finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
        }
      }
    } else {
      int pos = 0, gp = -1, gpEnd = 0, curId = 0, smallest = 0;
      double ilo = -INF, ihi = INF, dlb = INF;
      for (;;) {
        // fast-forward over the body pairs the reference would skip (body-pair-collision.cc:51-63)
        if (work && gp < 0) {
          while (pos < NP) {
            const int id = order[pos * SS];
            if (ivvalid[id * SS]) {  // valid_ on the whole path: interval = timeRange
              if (ihi > L) smallest = pos;
              ilo = (ilo < 0.0) ? 0.0 : ilo;
              ihi = (L < ihi) ? L : ihi;
              ++pos;
              continue;
            }
            const size_t o = ((size_t)id * S.K) * SS;
            if (ivContains(ivlo + o, ivhi + o, SS, ivcnt[id * SS], ilo, ihi)) { ++pos; continue; }
            curId = id;
            gp = T.bodypairs[id].pair_begin;
            gpEnd = T.bodypairs[id].pair_end;
            dlb = INF;
            break;
          }
        }
        const bool pending = work && gp >= 0;
        if (!__any_sync(FULL, pending)) break;
        if (pending) {
          // one narrow-phase test per lane, warp in step
          const DevPair pr = T.bgpairs[gp];
          const double d = pairLowerBound(T, pr, store, NS, mv.vertices, cnt);
          ++tests;
          if (d <= HPPB200_COLLISION_THRESHOLD) {
            finishEdge(false, HPPB200_REASON_COLLISION, t, DICH ? dichLast() : lastValid, gp);
            work = false;
          } else {
            if (d < dlb) dlb = d;
            ++gp;
            if (gp >= gpEnd) {
              // BodyPairCollision::validateConfiguration :72-96
              const DevBodyPair& B = T.bodypairs[curId];
              double hd, ht;
              if (dlb == INF) { hd = INF; ht = 0; }
              else {
                double Vm;
                if (SPLINE) hd = splineCollisionFreeInterval(T, B, Pp, mv.nq, L, t, dlb, Vm);
                else { Vm = maximalVelocity(B); hd = dlb / Vm; }
                ht = (Vm != 0) ? 2 * A.param / Vm : INF;
              }
              const double a = t - (hd + ht), b = t + (hd + ht);
              const size_t o = ((size_t)curId * S.K) * SS;
              int c = ivcnt[curId * SS];
              if (!ivUnion(ivlo + o, ivhi + o, SS, c, S.K, a, b)) {
                if (A.overflow_reason == HPPB200_REASON_ITER_CAP) ++caps; else ++overflows;
                finishEdge(false, A.overflow_reason, t, DICH ? dichLast() : lastValid, -1);
                work = false;
              } else {
                ivcnt[curId * SS] = (uint8_t)c;
                if (ivContains(ivlo + o, ivhi + o, SS, c, 0.0, L)) ivvalid[curId * SS] = 1;
                // validateIntervals: fold into the running intersection
                if (ihi > b) smallest = pos;
                ilo = (ilo < a) ? a : ilo;
                ihi = (b < ihi) ? b : ihi;
                ++pos;
                gp = -1;
              }
            }
          }
        }
      }
      // ================= D. end of a collision-free sample =========================================
      if (work) {
        // src/continuous-validation.cc:144-146: smallest interval first
        if (NP > 1 && smallest != 0) {
          const uint16_t a0 = order[0], as = order[smallest * SS];
          order[0] = as;
          order[smallest * SS] = a0;
        }
        if (PROG) {
          const double tprev = t;
          t = reverse ? ilo : ihi;
          lastValid = tprev;
          if (reverse ? t <= 0.0 : t >= L) { t = reverse ? 0.0 : L; finished++; }
          if (finished >= 2) finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
        } else {
          if (!ivUnion(sublo, subhi, NS, subCnt, S.KS, ilo, ihi)) {
            if (A.overflow_reason == HPPB200_REASON_ITER_CAP) ++caps; else ++overflows;
            finishEdge(false, A.overflow_reason, t, dichLast(), -1);
          } else if (ivContains(sublo, subhi, NS, subCnt, 0.0, L)) {
            finishEdge(true, HPPB200_REASON_VALID, CUDART_NAN, reverse ? 0.0 : L, -1);
          } else {
            // middle of the first non-validated gap (dichotomy.cc:103-109)
            double t0, t1;
            if (!reverse) { t0 = subhi[0]; t1 = subCnt > 1 ? sublo[NS] : L; }
            else { t0 = sublo[(size_t)(subCnt - 1) * NS]; t1 = subCnt > 1 ? subhi[(size_t)(subCnt - 2) * NS] : 0.0; }
            t = .5 * (t0 + t1);
          }
        }
      }
      // ---- mode 1: first sample done; park the edge for the resume pass ----
      if (A.mode == 1) {
        const bool cont = active;
        const unsigned m = __ballot_sync(FULL, cont);
        if (m) {
          const int leader = __ffs(m) - 1;
          unsigned long long b = 0;
          if (lane == leader) b = atomicAdd(A.cont_count, (unsigned long long)__popc(m));
          b = __shfl_sync(FULL, b, leader);
          if (cont) {
            const size_t NT = A.n_total;
            A.cont_list[b + __popc(m & ((1u << lane) - 1u))] = (uint32_t)e;
            A.sv_dbl[e] = t;
            A.sv_dbl[NT + e] = lastValid;
            A.sv_flags[e] = (uint8_t)finished;
            for (int id = 0; id < NP; ++id) {
              A.sv_order[(size_t)id * NT + e] = order[id * SS];
              A.sv_flags[(size_t)(1 + id) * NT + e] = ivcnt[id * SS];
              A.sv_flags[(size_t)(1 + NP + id) * NT + e] = ivvalid[id * SS];
              const size_t o = ((size_t)id * S.K) * SS;
              A.sv_dbl[(size_t)(2 + id) * NT + e] = ivlo[o];
              A.sv_dbl[(size_t)(2 + NP + id) * NT + e] = ivhi[o];
            }
            if (DICH) { A.sv_dbl[(size_t)(2 + 2 * NP) * NT + e] = sublo[0]; A.sv_dbl[(size_t)(3 + 2 * NP) * NT + e] = subhi[0]; }
            active = false;
          }
        }
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
    double d;
    if (J.type == HPPB200_JOINT_FREEFLYER) d = ffSquaredDistance(a + J.idx_q, b + J.idx_q, w2);
    else {
      const double dq = b[J.idx_q] - a[J.idx_q];
      d = (J.type <= HPPB200_JOINT_REVOLUTE_Z) ? w2 * (dq * dq) : dq * dq;
    }
    res += d;
  }
  L[i] = sqrt(res);
}

// Path::eval -> StraightPath::impl_compute (reference src/straight-path.cc:91-110), one thread per edge
__global__ void pathEvalKernel(ModelView mv, const double* __restrict__ q1, const double* __restrict__ q2,
                               const double* __restrict__ L, const double* __restrict__ t, size_t n, double* __restrict__ q) {
  const Tables T = makeTables(mv.blob, mv);
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const EdgeQ eq = edgeQ(q1 + i * (size_t)mv.nq, q2 + i * (size_t)mv.nq, t[i], L[i]);
  double* out = q + i * (size_t)mv.nq;
  for (int j = 0; j < mv.njoints; ++j) {
    const DevJoint& J = T.joints[j];
    if (J.type == HPPB200_JOINT_FREEFLYER) {
      double f[7];
      eq.freeflyer(J.idx_q, f);
      for (int k = 0; k < 7; ++k) out[J.idx_q + k] = f[k];
    } else out[J.idx_q] = eq(J.idx_q);
  }
}

void launchPathEval(Workspace* ws, const double* d_q1, const double* d_q2, const double* d_L, const double* d_t, size_t n, double* d_q) {
  if (n == 0) return;
  pathEvalKernel<<<(unsigned)((n + 255) / 256), 256, 0, ws->stream>>>(ws->view, d_q1, d_q2, d_L, d_t, n, d_q);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
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

template <int METHOD, bool SPLINE = false>
static void launchEdgeWalker(Workspace* ws, const EdgeArgs& A0, int K, int KS, size_t maxSlots) {
  const bool tree = modelIsTree(ws->model);
  auto kern = tree ? edgeWalkerKernel<METHOD, true, SPLINE> : edgeWalkerKernel<METHOD, false, SPLINE>;
  const ModelView& mv = ws->view;
  size_t smemBytes = 16 + mv.blob_bytes;
  const size_t nbp0 = METHOD == HPPB200_METHOD_DISCRETIZED ? 0 : (size_t)mv.nbodypairs;
  const size_t stateBytes = nbp0 * EDGE_THREADS * ((size_t)K * 16 + 4);
  const bool smemState = nbp0 > 0 && smemBytes + stateBytes <= 200u * 1024u;
  if (smemState) smemBytes += stateBytes;
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
  const size_t nbp = (METHOD == HPPB200_METHOD_DISCRETIZED || smemState) ? 0 : (size_t)mv.nbodypairs;
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
  S.smem_state = smemState ? 1 : 0;
  EdgeArgs A = A0;
  A.counter = (unsigned long long*)(base + o_counter);

  HPPB200_CUDA_CHECK(cudaMemsetAsync(A.counter, 0, sizeof(unsigned long long), ws->stream));
  kern<<<(unsigned)grid, EDGE_THREADS, smemBytes, ws->stream>>>(mv, A, S, ws->d_stats);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
}

void launchValidateEdgesCore(Workspace* ws, int method, double param, int reverse, const double* d_q1, const double* d_q2,
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
  const size_t allSlots = ~size_t(0);
  if (method == HPPB200_METHOD_DISCRETIZED) { launchEdgeWalker<HPPB200_METHOD_DISCRETIZED>(ws, A, 1, 1, allSlots); return; }  // (no lists, no overflow)
  DeviceStats before{}, after;
  // (only the overflow counter is compared; it changes in edge calls only, and every edge call ends by reading it: a one-item
  //  call does not pay a copy and a wait for a number the host already has)
  if (ws->overflow_known) before.overflow = ws->overflow_cached;
  else HPPB200_CUDA_CHECK(cudaMemcpyAsync(&before, ws->d_stats, sizeof(before), cudaMemcpyDeviceToHost, ws->stream));
  ws->overflow_known = false;
  // Progressive on models whose per-edge state fits in shared memory: the control/narrow-phase
  // engine of kernels_continuous.cu.  Anything else: the generic walker below.
  bool engineDone = false;
  // Large batches: round-synchronous wavefront engine (kernels_wavefront.cu); its ~600 rounds of a few kernels
  // each are a fixed latency that small batches do not amortise, those go to the CTA engine.
  // (Dichotomy: the CTA-per-edge kernel below takes the batches under 50 k edges and the wavefront's tail;
  //  measured 200 k UR5 edges: wavefront 1.11 M edges/s, CTA per edge 0.64 M, walker 0.73 M; 50 k: 0.65 M vs 0.5 M)
  const size_t wfMin = method == HPPB200_METHOD_DICHOTOMY ? ws->wavefront_min / 2 : ws->wavefront_min;
  if (method != HPPB200_METHOD_DISCRETIZED && ws->use_engine && ws->engine_kind >= 2 && n >= wfMin)
    engineDone = launchWavefront(ws, method, param, reverse, d_q1, d_q2, d_L, n, d_valid, d_tfail, d_tlast, d_pair,
                                            d_reason, REASON_OVERFLOW_INTERNAL);
  // Progressive, small batches: one CTA per edge (kernels_tail.cu), the latency of a sample instead of the CTA engine's ticks
  // (models with many body pairs: the CTA-per-edge kernel also wins on mid-sized batches, e.g. what the first-sample screen leaves of
  //  BASELINE configs[3]: 87 k edges of the 273-pair humanoid 59.9 ms against 64.2 ms through the CTA engine; 1 M edges 169 vs 163 ms)
  const size_t progTailMax = ws->view.nbodypairs > 128 ? std::max<size_t>(ws->prog_tail_max, ws->prog_tail_max ? 131072 : 0) : ws->prog_tail_max;
  if (!engineDone && method == HPPB200_METHOD_PROGRESSIVE && ws->use_tail && n <= progTailMax) {
    const int k = std::min(64, (int)(65536 / (16 * std::max(1, (int)ws->view.nbodypairs))));
    if (k >= 8)
      engineDone = launchEdgeTail(ws, method, param, reverse, d_q1, d_q2, d_L, nullptr, n, d_valid, d_tfail, d_tlast, d_pair, d_reason,
                                  REASON_OVERFLOW_INTERNAL, true, nullptr);
  }
  if (!engineDone && method == HPPB200_METHOD_PROGRESSIVE && ws->use_engine)
    engineDone = launchProgressiveEngine(ws, param, reverse, d_q1, d_q2, d_L, n, d_valid, d_tfail, d_tlast, d_pair, d_reason,
                                         REASON_OVERFLOW_INTERNAL);
  // Dichotomy, one CTA per edge (kernels_tail.cu): level 1 keeps the per-pair interval lists in shared memory,
  // level 2 (64-entry lists in global memory) takes what outgrows level 1; small batches start there directly
  const int smemK = std::min(64, (int)(65536 / (16 * std::max(1, (int)ws->view.nbodypairs))));
  auto tailPass = [&](const uint32_t* list, size_t cnt, int level) {
    if (level == 1 && (!ws->tail_smem || smemK < 8)) return false;
    return launchDichotomyTail(ws, param, reverse, d_q1, d_q2, d_L, list, cnt, d_valid, d_tfail, d_tlast, d_pair, d_reason,
                               REASON_OVERFLOW_INTERNAL, level == 1);
  };
  int tailLevel = 0;
  if (!engineDone && method == HPPB200_METHOD_DICHOTOMY && ws->use_tail && n <= ws->tail_max) {
    for (int level = 1; level <= 2 && !engineDone; ++level)
      if (tailPass(nullptr, n, level)) { engineDone = true; tailLevel = level; }
  }
  if (!engineDone) {
  // pass 1: first sample of every edge (uniform warps); pass 2: resume the survivors
  const size_t NPb = (size_t)ws->view.nbodypairs;
  ws->d_sv_dbl.reserve((2 + 2 * NPb + 2) * n);
  ws->d_sv_order.reserve(NPb * n);
  ws->d_sv_flags.reserve((1 + 2 * NPb) * n);
  ws->d_cont.reserve(n);
  ws->d_count.reserve(2);
  A.sv_dbl = ws->d_sv_dbl.p; A.sv_order = ws->d_sv_order.p; A.sv_flags = ws->d_sv_flags.p;
  A.cont_list = ws->d_cont.p; A.cont_count = ws->d_count.p + 1; A.n_total = n;
  HPPB200_CUDA_CHECK(cudaMemsetAsync(A.cont_count, 0, sizeof(unsigned long long), ws->stream));
  A.mode = 1;
  if (method == HPPB200_METHOD_PROGRESSIVE) launchEdgeWalker<HPPB200_METHOD_PROGRESSIVE>(ws, A, 1, 1, allSlots);
  else launchEdgeWalker<HPPB200_METHOD_DICHOTOMY>(ws, A, 1, 1, allSlots);
  unsigned long long ncont = 0;
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&ncont, A.cont_count, sizeof(ncont), cudaMemcpyDeviceToHost, ws->stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  if (ncont) {
    EdgeArgs B = A;
    B.mode = 2;
    B.edge_list = ws->d_cont.p;
    B.n = (size_t)ncont;
    if (method == HPPB200_METHOD_PROGRESSIVE) launchEdgeWalker<HPPB200_METHOD_PROGRESSIVE>(ws, B, 1, 1, allSlots);
    else launchEdgeWalker<HPPB200_METHOD_DICHOTOMY>(ws, B, 12, 96, allSlots);
  }
  }
  A.mode = 0;
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&after, ws->d_stats, sizeof(after), cudaMemcpyDeviceToHost, ws->stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  const unsigned long long nover = after.overflow - before.overflow;
  if (std::getenv("HPPB200_WF_TRACE")) std::fprintf(stderr, "[edges] %llu of %zu edges go to the large-list walker pass\n", nover, n);
  if (nover == 0) { ws->overflow_known = true; ws->overflow_cached = after.overflow; return; }
  // second pass: gather the overflowed edges and walk them with large interval lists
  ws->d_list.reserve(nover + 1);
  ws->d_count.reserve(1);
  HPPB200_CUDA_CHECK(cudaMemsetAsync(ws->d_count.p, 0, sizeof(unsigned long long), ws->stream));
  collectOverflowKernel<<<(unsigned)((n + 255) / 256), 256, 0, ws->stream>>>(d_reason, n, ws->d_list.p, ws->d_count.p);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  A.edge_list = ws->d_list.p;
  A.n = (size_t)nover;
  if (method == HPPB200_METHOD_DICHOTOMY && ws->use_tail) {
    unsigned long long cur = nover;
    DeviceStats last = after;
    for (int level = tailLevel + 1; level <= 2; ++level) {
      if (level == 2 && tailLevel == 1 && smemK >= 64) continue;  // same capacity as the pass that overflowed
      if (!tailPass(ws->d_list.p, (size_t)cur, level)) continue;
      tailLevel = level;
      DeviceStats now;
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(&now, ws->d_stats, sizeof(now), cudaMemcpyDeviceToHost, ws->stream));
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
      const unsigned long long left = now.overflow - last.overflow;
      last = now;
      if (std::getenv("HPPB200_WF_TRACE")) std::fprintf(stderr, "[edges] CTA-per-edge pass %d: %llu of %llu edges go on\n", level, left, cur);
      if (left == 0) return;
      HPPB200_CUDA_CHECK(cudaMemsetAsync(ws->d_count.p, 0, sizeof(unsigned long long), ws->stream));
      collectOverflowKernel<<<(unsigned)((n + 255) / 256), 256, 0, ws->stream>>>(d_reason, n, ws->d_list.p, ws->d_count.p);
      HPPB200_CUDA_CHECK(cudaGetLastError());
      ws->stats.kernel_launches += 1;
      cur = left;
    }
    A.n = (size_t)cur;
  }
  A.overflow_reason = HPPB200_REASON_ITER_CAP;
  const size_t slots2 = 2048;  // 255-entry lists: ~4 KB per body pair per slot
  if (method == HPPB200_METHOD_PROGRESSIVE) launchEdgeWalker<HPPB200_METHOD_PROGRESSIVE>(ws, A, 255, 1, slots2);
  else launchEdgeWalker<HPPB200_METHOD_DICHOTOMY>(ws, A, 255, 2048, slots2);
}

// Edges that are path::Spline<BernsteinBasis, 3> (reference tests/test-continuous-validation.cc:288-427):
// the lane-per-edge walker with the spline's eval / velocityBound and the two-round refinement of BodyPairCollision.
void launchValidateSplineEdges(Workspace* ws, int method, double param, int reverse, const double* d_base, const double* d_params,
                               const double* d_L, size_t n, uint32_t flags, uint8_t* d_valid, double* d_tfail, double* d_tlast,
                               int32_t* d_pair, uint8_t* d_reason) {
  if (n == 0) return;
  if (n > 0xffffffffull) throw std::invalid_argument("at most 2^32-1 edges per call");
  if (ws->view.nbodypairs > 65535) throw std::runtime_error("too many body pairs");
  if (ws->model->nq != ws->model->nv) throw std::invalid_argument("spline paths: vector-space configuration spaces only (no free-flyer joints)");
  if (!d_reason) {
    ws->d_reason_int.reserve(n);
    d_reason = ws->d_reason_int.p;
  }
  EdgeArgs A{};
  A.q1 = d_base; A.q2 = nullptr; A.P = d_params; A.L = d_L; A.n = n; A.flags = flags; A.reverse = reverse; A.param = param;
  A.max_iter = ws->max_iterations;
  A.valid = d_valid; A.t_fail = d_tfail; A.t_last = d_tlast; A.pair = d_pair; A.reason = d_reason;
  A.overflow_reason = REASON_OVERFLOW_INTERNAL;
  const size_t allSlots = ~size_t(0);
  if (method == HPPB200_METHOD_DISCRETIZED) { launchEdgeWalker<HPPB200_METHOD_DISCRETIZED, true>(ws, A, 1, 1, allSlots); return; }
  DeviceStats before{}, after;
  // (only the overflow counter is compared; it changes in edge calls only, and every edge call ends by reading it: a one-item
  //  call does not pay a copy and a wait for a number the host already has)
  if (ws->overflow_known) before.overflow = ws->overflow_cached;
  else HPPB200_CUDA_CHECK(cudaMemcpyAsync(&before, ws->d_stats, sizeof(before), cudaMemcpyDeviceToHost, ws->stream));
  ws->overflow_known = false;
  // (Progressive keeps one connected validated interval per body pair whatever the path: the next sample lies inside
  //  every pair's last interval.  Dichotomy: short lists, the edges that outgrow them are redone with long ones.)
  if (method == HPPB200_METHOD_PROGRESSIVE) launchEdgeWalker<HPPB200_METHOD_PROGRESSIVE, true>(ws, A, 1, 1, allSlots);
  else launchEdgeWalker<HPPB200_METHOD_DICHOTOMY, true>(ws, A, 12, 96, allSlots);
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&after, ws->d_stats, sizeof(after), cudaMemcpyDeviceToHost, ws->stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  const unsigned long long nover = after.overflow - before.overflow;
  if (nover == 0) return;
  ws->d_list.reserve(nover + 1);
  ws->d_count.reserve(1);
  HPPB200_CUDA_CHECK(cudaMemsetAsync(ws->d_count.p, 0, sizeof(unsigned long long), ws->stream));
  collectOverflowKernel<<<(unsigned)((n + 255) / 256), 256, 0, ws->stream>>>(d_reason, n, ws->d_list.p, ws->d_count.p);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
  A.edge_list = ws->d_list.p;
  A.n = (size_t)nover;
  A.overflow_reason = HPPB200_REASON_ITER_CAP;
  const size_t slots2 = 2048;
  if (method == HPPB200_METHOD_PROGRESSIVE) launchEdgeWalker<HPPB200_METHOD_PROGRESSIVE, true>(ws, A, 255, 1, slots2);
  else launchEdgeWalker<HPPB200_METHOD_DICHOTOMY, true>(ws, A, 255, 2048, slots2);
}

}  // namespace hppb200
