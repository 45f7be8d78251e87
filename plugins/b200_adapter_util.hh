// Pieces of the hpp-core adapter (plugins/hpp_b200_plugin.cc) that do not depend on hpp-core's headers, so that
// they can be compiled and tested in the development image with mock node / component classes
// (tests/plugin_mock/test_adapter_util.cc, run by tests/test_plugin_static.py).
#pragma once
#include <algorithm>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace hpp {
namespace b200 {

// Connected-component labels of the device mirror of the roadmap nodes.
//
// The reference never tells the NearestNeighbor object about merges: Roadmap::merge (src/roadmap.cc:298-310) only
// calls cc1->merge(cc2) and nearestNeighbor::Basic scans connectedComponent->nodes() live
// (src/nearest-neighbor/basic.cc:52-118).  So the labels are DERIVED from the roadmap: before a query restricted to a
// component, the component's live node list is compared with what the mirror holds under its label; nodes that
// arrived through merges are relabelled (whole old components at a time: a merge moves all of them).
// Components are keyed by address AND by a weak reference: a freed component whose address is reused gets a new label.
//
// CC: component type with `nodes()` (iterable of Node*) ; CCPtr = std::shared_ptr<CC>-like with get() and conversion to
// the weak pointer type WK (lock() returns CCPtr).
template <class CC, class CCPtr, class WK, class NodeT>
class ComponentLabels {
 public:
  typedef std::function<void(int32_t from, int32_t to)> Relabel;  // hppb200_knn_relabel on the device mirror
  // a node of a live component that the mirror was never given (a roadmap restored by serialisation,
  // src/serialization.cc:86-99): the callback appends it to the device mirror with `label`; it is then registered here
  typedef std::function<void(NodeT* node, int32_t label)> Adopt;
  explicit ComponentLabels(Relabel r, Adopt a = Adopt()) : relabel_(std::move(r)), adopt_(std::move(a)) {}

  void clear() { byPtr_.clear(); nodeLabel_.clear(); index_.clear(); count_.clear(); next_ = 0; }
  std::size_t size() const { return nodeLabel_.size(); }

  // label of a component (a new one if the pointer is unknown or was recycled)
  int32_t label(const CCPtr& cc) {
    auto it = byPtr_.find(cc.get());
    if (it != byPtr_.end()) {
      CCPtr alive = it->second.first.lock();
      if (alive && alive.get() == cc.get()) return it->second.second;
      byPtr_.erase(it);  // the component that had this address is gone
    }
    const int32_t l = next_++;
    byPtr_.insert({cc.get(), {WK(cc), l}});
    count_[l] = 0;
    return l;
  }

  // a node was appended to the mirror (index = insertion order); returns the label to store with it
  int32_t addNode(NodeT* node, const CCPtr& cc) {
    const int32_t l = label(cc);
    index_[node] = nodeLabel_.size();
    nodeLabel_.push_back(l);
    ++count_[l];
    return l;
  }

  // call before every query restricted to `cc`; returns its label with the mirror's labels brought up to date
  int32_t sync(const CCPtr& cc) {
    const int32_t l = label(cc);
    const std::size_t live = cc->nodes().size();
    if (live == count_[l]) return l;
    std::vector<int32_t> olds;
    for (NodeT* n : cc->nodes()) {
      auto it = index_.find(n);
      if (it == index_.end()) {
        if (!adopt_) continue;
        adopt_(n, l);
        index_[n] = nodeLabel_.size();
        nodeLabel_.push_back(l);
        ++count_[l];
        continue;
      }
      const int32_t o = nodeLabel_[it->second];
      if (o != l && std::find(olds.begin(), olds.end(), o) == olds.end()) olds.push_back(o);
    }
    for (int32_t o : olds) {
      relabel_(o, l);
      for (int32_t& x : nodeLabel_)
        if (x == o) x = l;
      count_[l] += count_[o];
      count_[o] = 0;
    }
    return l;
  }

  int32_t labelOfNode(std::size_t i) const { return nodeLabel_[i]; }

 private:
  Relabel relabel_;
  Adopt adopt_;
  std::map<const CC*, std::pair<WK, int32_t>> byPtr_;
  std::vector<int32_t> nodeLabel_;                 // host snapshot of the mirror's labels
  std::unordered_map<NodeT*, std::size_t> index_;  // node -> mirror index
  std::map<int32_t, std::size_t> count_;           // nodes per label in the mirror
  int32_t next_ = 0;
};

// Pool of engine workspaces (one CUDA stream + scratch each), owned by the scene: borrowed for the duration of one
// validate call, like the reference's pool of BodyPairCollision copies (src/continuous-validation.cc:193-205).
// Destroyed (with all its workspaces) BEFORE the model it was created for.
template <class WS>
class WorkspacePool {
 public:
  typedef std::function<WS*()> Create;
  typedef std::function<void(WS*)> Destroy;
  WorkspacePool(Create c, Destroy d) : create_(std::move(c)), destroy_(std::move(d)) {}
  ~WorkspacePool() { clear(); }
  WorkspacePool(const WorkspacePool&) = delete;
  WorkspacePool& operator=(const WorkspacePool&) = delete;

  class Lease {
   public:
    Lease(WorkspacePool* p, WS* w) : pool_(p), ws_(w) {}
    Lease(Lease&& o) noexcept : pool_(o.pool_), ws_(o.ws_) { o.ws_ = nullptr; }
    Lease(const Lease&) = delete;
    ~Lease() { if (ws_) pool_->release(ws_); }
    WS* get() const { return ws_; }
   private:
    WorkspacePool* pool_;
    WS* ws_;
  };

  Lease acquire() {
    {
      std::lock_guard<std::mutex> lock(mu_);
      if (!free_.empty()) { WS* w = free_.back(); free_.pop_back(); ++out_; return Lease(this, w); }
      ++out_;
    }
    WS* w = nullptr;
    try { w = create_(); } catch (...) { std::lock_guard<std::mutex> lock(mu_); --out_; throw; }
    return Lease(this, w);
  }
  // all workspaces must have been returned (the scene rebuilds under its own lock, with no call in flight)
  void clear() {
    std::lock_guard<std::mutex> lock(mu_);
    for (WS* w : free_) destroy_(w);
    free_.clear();
  }
  std::size_t idle() const { std::lock_guard<std::mutex> lock(mu_); return free_.size(); }
  std::size_t borrowed() const { std::lock_guard<std::mutex> lock(mu_); return out_; }

 private:
  void release(WS* w) { std::lock_guard<std::mutex> lock(mu_); free_.push_back(w); --out_; }
  Create create_;
  Destroy destroy_;
  mutable std::mutex mu_;
  std::vector<WS*> free_;
  std::size_t out_ = 0;
};

}  // namespace b200
}  // namespace hpp
