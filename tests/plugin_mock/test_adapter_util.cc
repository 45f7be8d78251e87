// CPU mock test of plugins/b200_adapter_util.hh: drives ComponentLabels through the sequence a Roadmap produces
// (addNode into a fresh component, addEdge -> ConnectedComponent::merge WITHOUT any call to NearestNeighbor::merge,
// reference src/roadmap.cc:298-310), including a component freed and another allocated at the same address, and
// checks that component-restricted queries see exactly the component's live node list.  Also WorkspacePool.
#include <cassert>
#include <cstdio>
#include <cstdlib>
#include <new>
#include <random>
#include <set>
#include <thread>

#include "../../plugins/b200_adapter_util.hh"

struct Node;
struct CC {
  std::vector<Node*> nodes_;
  const std::vector<Node*>& nodes() const { return nodes_; }
};
typedef std::shared_ptr<CC> CCPtr;
struct Node {
  CCPtr cc;
};

struct Mirror {  // stands for the device mirror: label per node in insertion order
  std::vector<int32_t> label;
  void relabel(int32_t from, int32_t to) {
    for (int32_t& x : label)
      if (x == from) x = to;
  }
  std::set<std::size_t> withLabel(int32_t l) const {
    std::set<std::size_t> s;
    for (std::size_t i = 0; i < label.size(); ++i)
      if (label[i] == l) s.insert(i);
    return s;
  }
};

typedef hpp::b200::ComponentLabels<CC, CCPtr, std::weak_ptr<CC>, Node> Labels;

static void checkQuery(Labels& L, Mirror& M, const CCPtr& cc, const std::vector<Node*>& all) {
  const int32_t l = L.sync(cc);
  std::set<std::size_t> want;
  for (Node* n : cc->nodes())
    for (std::size_t i = 0; i < all.size(); ++i)
      if (all[i] == n) want.insert(i);
  if (M.withLabel(l) != want) { std::fprintf(stderr, "component query sees the wrong node set\n"); std::exit(1); }
}

int main() {
  Mirror M;
  Labels L([&](int32_t a, int32_t b) { M.relabel(a, b); });
  std::vector<Node*> all;
  std::vector<CCPtr> comps;
  std::mt19937 rng(7);
  auto addNode = [&]() {  // Roadmap::addNode: new node in its own new component, then nearestNeighbor_->addNode
    CCPtr cc(new CC);
    Node* n = new Node{cc};
    cc->nodes_.push_back(n);
    comps.push_back(cc);
    all.push_back(n);
    M.label.push_back(L.addNode(n, cc));
  };
  auto merge = [&](std::size_t i, std::size_t j) {  // ConnectedComponent::merge (src/connected-component.cc:43-53)
    CCPtr a = comps[i], b = comps[j];
    for (Node* n : b->nodes_) { n->cc = a; a->nodes_.push_back(n); }
    b->nodes_.clear();
    comps.erase(comps.begin() + j);  // Roadmap erases the absorbed component; its last reference dies here
  };
  for (int step = 0; step < 3000; ++step) {
    const int r = rng() % 10;
    if (r < 5 || comps.size() < 2) addNode();
    else if (r < 8) {
      std::size_t i = rng() % comps.size(), j = rng() % comps.size();
      if (i != j) merge(i, j);
    }
    if (!comps.empty()) checkQuery(L, M, comps[rng() % comps.size()], all);
  }
  for (const CCPtr& cc : comps) checkQuery(L, M, cc, all);

  // a component freed and a new one constructed at the SAME address must not inherit the old label
  {
    alignas(CC) static unsigned char buf[sizeof(CC)];
    auto make = [&]() { return CCPtr(new (buf) CC, [](CC* p) { p->~CC(); }); };
    CCPtr first = make();
    Node* n1 = new Node{first};
    first->nodes_.push_back(n1);
    all.push_back(n1);
    M.label.push_back(L.addNode(n1, first));
    const int32_t l1 = L.sync(first);
    // merged away into comps[0] (no query on comps[0] in between), then destroyed
    n1->cc = comps[0]; comps[0]->nodes_.push_back(n1); first->nodes_.clear();
    first.reset();
    CCPtr second = make();
    assert((void*)second.get() == (void*)buf);
    Node* n2 = new Node{second};
    second->nodes_.push_back(n2);
    all.push_back(n2);
    M.label.push_back(L.addNode(n2, second));
    const int32_t l2 = L.sync(second);
    if (l2 == l1) { std::fprintf(stderr, "recycled component address inherited a stale label\n"); return 1; }
    checkQuery(L, M, second, all);
    checkQuery(L, M, comps[0], all);
  }

  // a roadmap the mirror never saw (deserialised): nodes are adopted at the first component query
  {
    Mirror M2;
    std::vector<Node*> all2;
    Labels L2([&](int32_t a, int32_t b) { M2.relabel(a, b); }, [&](Node* n, int32_t l) { all2.push_back(n); M2.label.push_back(l); });
    for (const CCPtr& cc : comps) checkQuery(L2, M2, cc, all2);
    std::size_t total = 0;
    for (const CCPtr& cc : comps) total += cc->nodes().size();
    if (all2.size() != total || L2.size() != total) { std::fprintf(stderr, "adoption missed nodes\n"); return 1; }
    for (const CCPtr& cc : comps) checkQuery(L2, M2, cc, all2);
  }

  // WorkspacePool: concurrent borrow / return, every workspace destroyed exactly once
  {
    int created = 0, destroyed = 0;
    std::mutex mu;
    {
      hpp::b200::WorkspacePool<int> pool([&]() { std::lock_guard<std::mutex> l(mu); ++created; return new int(0); },
                                         [&](int* p) { std::lock_guard<std::mutex> l(mu); ++destroyed; delete p; });
      std::vector<std::thread> th;
      for (int t = 0; t < 8; ++t)
        th.emplace_back([&]() {
          for (int i = 0; i < 2000; ++i) {
            auto lease = pool.acquire();
            int* w = lease.get();
            if (*w != 0) { std::fprintf(stderr, "workspace handed to two threads\n"); std::exit(1); }
            *w = 1; *w = 0;
          }
        });
      for (auto& t : th) t.join();
      if (pool.borrowed() != 0 || pool.idle() != (std::size_t)created || created > 8) { std::fprintf(stderr, "pool accounting\n"); return 1; }
    }
    if (created != destroyed) { std::fprintf(stderr, "pool leaked workspaces\n"); return 1; }
  }
  std::printf("adapter util ok: %zu nodes, %zu components\n", all.size(), comps.size());
  return 0;
}
