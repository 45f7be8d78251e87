// Host side of the batched kPRM* front-end: the O(N k) array passes between the engine's three
// batched calls, in C++ (round 1 did them in numpy: 11.5 s of the 23.5 s of a 1 M-node roadmap).
//
//  hppb200_kprm_link_candidates  kPrmStar::connectNodeToClosestNeighbors (reference src/path-planner/k-prm-star.cc:153-193)
//      walks node i's k nearest neighbours nearest first and skips a neighbour j that already has an edge j -> i.  With
//      every node's neighbour list in hand, a directed candidate (i, j) is either validated whatever happens (`first`),
//      or only if the mirrored candidate (j, i), met earlier because j < i and i is among j's neighbours, was found
//      invalid (`second`; mirror = the flat slot of (j, i) in j's list, where its outcome will be).
//  hppb200_roadmap_merge_edges   Roadmap::addEdges + ConnectedComponent::merge (reference src/roadmap.cc:209-255) for
//      a list of edges in insertion order: the component of the edge's second node joins the first node's; labels out.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../../include/hppb200.h"

namespace hppb200 {
void setLastError(const std::string& s);  // capi.cu
}

namespace {
template <class F>
void parallelFor(size_t n, F f) {
  unsigned nt = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
  if (n < 65536) nt = 1;
  if (nt == 1) { f(0, n); return; }
  std::vector<std::thread> th;
  const size_t per = (n + nt - 1) / nt;
  for (unsigned t = 0; t < nt; ++t) {
    const size_t lo = std::min(n, t * per), hi = std::min(n, lo + per);
    if (lo < hi) th.emplace_back([=] { f(lo, hi); });
  }
  for (auto& x : th) x.join();
}
}  // namespace

extern "C" {

int hppb200_kprm_link_candidates(size_t n, int k, const int64_t* nbrs, const double* dist, int symmetric_distance,
                                 uint8_t* kind, int64_t* mirror) {
  if (!nbrs || !kind || !mirror || k <= 0 || (symmetric_distance && !dist)) {
    hppb200::setLastError("hppb200_kprm_link_candidates: null argument or k <= 0");
    return HPPB200_ERR_INVALID_ARGUMENT;
  }
  const size_t K = (size_t)k;
  parallelFor(n, [=](size_t lo, size_t hi) {
    for (size_t i = lo; i < hi; ++i) {
      for (size_t s = 0; s < K; ++s) {
        const size_t slot = i * K + s;
        const int64_t j = nbrs[slot];
        kind[slot] = 0;
        mirror[slot] = -1;
        if (j < 0 || (size_t)j == i || (size_t)j >= n) continue;
        kind[slot] = 1;
        if ((size_t)j > i) continue;  // j's own walk comes later: (i, j) is tried whatever happens
        // is i among j's neighbours, and where?
        const int64_t* lj = nbrs + (size_t)j * K;
        int64_t pos = -1;
        if (symmetric_distance) {
          // WeighedDistance on a vector space is symmetric bit for bit and a list is sorted by the key (distance, index):
          // binary search for the key (d(i, j), i) in j's list
          const double* dj = dist + (size_t)j * K;
          const double d = dist[slot];
          size_t a = 0, b = K;
          while (a < b) {
            const size_t m = (a + b) / 2;
            const bool less = lj[m] >= 0 && (dj[m] < d || (dj[m] == d && lj[m] < (int64_t)i));
            if (less) a = m + 1; else b = m;
          }
          if (a < K && lj[a] == (int64_t)i) pos = (int64_t)a;
        } else {
          for (size_t m = 0; m < K; ++m)
            if (lj[m] == (int64_t)i) { pos = (int64_t)m; break; }
        }
        if (pos >= 0) { kind[slot] = 2; mirror[slot] = (int64_t)((size_t)j * K + (size_t)pos); }
      }
    }
  });
  return HPPB200_OK;
}

int hppb200_roadmap_merge_edges(size_t n, int64_t* labels, const int64_t* pairs, size_t nedges) {
  if (!labels || (!pairs && nedges)) {
    hppb200::setLastError("hppb200_roadmap_merge_edges: null argument");
    return HPPB200_ERR_INVALID_ARGUMENT;
  }
  try {
    // union-find over the nodes; the root of a set carries the component's label.  Nodes that share a label on entry
    // are one component: link them first (no label changes).
    std::vector<int64_t> parent(n), lab(labels, labels + n);
    std::vector<uint32_t> size(n, 1u);
    for (size_t i = 0; i < n; ++i) parent[i] = (int64_t)i;
    auto root = [&](int64_t x) {
      while (parent[x] != x) { parent[x] = parent[parent[x]]; x = parent[x]; }
      return x;
    };
    auto unite = [&](int64_t ra, int64_t rb, int64_t label) {  // ra, rb roots, ra != rb
      if (size[ra] < size[rb]) std::swap(ra, rb);
      parent[rb] = ra;
      size[ra] += size[rb];
      lab[ra] = label;
    };
    {
      std::vector<size_t> order(n);
      for (size_t i = 0; i < n; ++i) order[i] = i;
      std::sort(order.begin(), order.end(), [&](size_t a, size_t b) { return labels[a] < labels[b] || (labels[a] == labels[b] && a < b); });
      for (size_t t = 1; t < n; ++t)
        if (labels[order[t]] == labels[order[t - 1]]) {
          const int64_t ra = root((int64_t)order[t - 1]), rb = root((int64_t)order[t]);
          if (ra != rb) unite(ra, rb, labels[order[t]]);
        }
    }
    for (size_t e = 0; e < nedges; ++e) {
      const int64_t a = pairs[2 * e], b = pairs[2 * e + 1];
      if (a < 0 || b < 0 || (size_t)a >= n || (size_t)b >= n) throw std::invalid_argument("hppb200_roadmap_merge_edges: node index out of range");
      const int64_t ra = root(a), rb = root(b);
      if (ra != rb) unite(ra, rb, lab[ra]);  // ConnectedComponent::merge: the second component joins the first
    }
    for (size_t i = 0; i < n; ++i) labels[i] = lab[root((int64_t)i)];
    return HPPB200_OK;
  } catch (const std::invalid_argument& e) {
    hppb200::setLastError(e.what());
    return HPPB200_ERR_INVALID_ARGUMENT;
  } catch (const std::exception& e) {
    hppb200::setLastError(e.what());
    return HPPB200_ERR_RUNTIME;
  }
}

}  // extern "C"
