// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_math.hpp header).
//
// nearestNeighbor::Basic restated (src/nearest-neighbor/basic.cc:52-181): linear scan
// in insertion order, strict '<' comparisons, size-K max-heap for k-NN.
// PARITY UNPINNED (tests/roadmap-1.cc:336-424 prints only).
#pragma once
#include <queue>

#include "oracle_validation.hpp"

namespace oracle {

struct Roadmap {
  const Device* dev;
  std::vector<double> nodes;  // [n][nq] in insertion (scan) order
  std::vector<int> cc;        // connected-component label per node
  explicit Roadmap(const Device& d) : dev(&d) {}
  size_t size() const { return cc.size(); }
  void addNode(const double* q, int label) {
    nodes.insert(nodes.end(), q, q + dev->nq);
    cc.push_back(label);
  }
};

typedef std::pair<double, long> DistAndNode;
struct DistAndNodeComp {
  bool operator()(const DistAndNode& a, const DistAndNode& b) const { return a.first < b.first; }
};

// basic.cc:52-73 — 1-NN, first-seen node wins ties
inline long nnSearch(const Roadmap& rm, const double* q, int cc, double& distance) {
  long result = -1;
  distance = kInf;
  for (size_t i = 0; i < rm.size(); ++i) {
    if (cc >= 0 && rm.cc[i] != cc) continue;
    double d = weighedDistance(*rm.dev, rm.nodes.data() + i * rm.dev->nq, q);
    if (d < distance) { distance = d; result = (long)i; }
  }
  return result;
}

// basic.cc:94-118 — K-NN; output nearest first; distance = K-th distance
inline std::vector<DistAndNode> nnKnearest(const Roadmap& rm, const double* q, int cc, size_t K, double& distance) {
  std::priority_queue<DistAndNode, std::vector<DistAndNode>, DistAndNodeComp> ns;
  distance = kInf;
  for (size_t i = 0; i < rm.size(); ++i) {
    if (cc >= 0 && rm.cc[i] != cc) continue;
    double d = weighedDistance(*rm.dev, rm.nodes.data() + i * rm.dev->nq, q);
    if (ns.size() < K) ns.push(DistAndNode(d, (long)i));
    else if (ns.top().first > d) { ns.pop(); ns.push(DistAndNode(d, (long)i)); }
  }
  std::vector<DistAndNode> out;
  if (!ns.empty()) distance = ns.top().first;
  while (!ns.empty()) { out.insert(out.begin(), ns.top()); ns.pop(); }
  return out;
}

// basic.cc:170-181 — strict d < maxDistance, scan order
inline std::vector<long> nnWithinBall(const Roadmap& rm, const double* q, int cc, double maxDistance) {
  std::vector<long> out;
  for (size_t i = 0; i < rm.size(); ++i) {
    if (cc >= 0 && rm.cc[i] != cc) continue;
    if (weighedDistance(*rm.dev, rm.nodes.data() + i * rm.dev->nq, q) < maxDistance) out.push_back((long)i);
  }
  return out;
}

}  // namespace oracle
