// Multi-GPU variant of the validation entry points: a device list, sharded internally.
//
// Configurations and edges are independent units (the reference exploits the same property with OpenMP and pooled
// state, tests/test-continuous-validation.cc:324-343), the model tables are small and replicated, so the path has no
// data-path collective in this form: one persistent host thread per device validates the contiguous index range
// [r n / N, (r + 1) n / N) of the caller's HOST arrays through the single-GPU host-pointer path (sliced copies
// overlapped with the kernels) and writes its results straight into the caller's output arrays: the "gather of the
// validity masks" is the D2H copy of every shard into its place.  (The one-process-per-GPU form with an NCCL
// all-gather of device-resident masks is bench.py / hpp-core_b200/sharding.py.)
#include <condition_variable>
#include <functional>
#include <memory>
#include <thread>

#include "engine.hpp"

namespace hppb200 {
void setLastError(const std::string& s);
}

namespace {

struct Worker {
  int device = 0;
  hppb200_workspace* ws = nullptr;
  std::thread thread;
  std::mutex mu;
  std::condition_variable cv;
  std::function<int()> task;  // returns an HPPB200_* status
  bool hasTask = false, done = false, quit = false;
  int rc = 0;
  std::string error;

  void loop() {
    cudaSetDevice(device);
    for (;;) {
      std::function<int()> t;
      {
        std::unique_lock<std::mutex> lock(mu);
        cv.wait(lock, [&] { return hasTask || quit; });
        if (quit) return;
        t = task;
        hasTask = false;
      }
      int r = t();
      std::string e = r == HPPB200_OK ? std::string() : std::string(hppb200_last_error());  // thread-local message of THIS thread
      {
        std::lock_guard<std::mutex> lock(mu);
        rc = r;
        error = e;
        done = true;
      }
      cv.notify_all();
    }
  }
  void submit(std::function<int()> t) {
    {
      std::lock_guard<std::mutex> lock(mu);
      task = std::move(t);
      hasTask = true;
      done = false;
    }
    cv.notify_all();
  }
  int wait() {
    std::unique_lock<std::mutex> lock(mu);
    cv.wait(lock, [&] { return done; });
    return rc;
  }
};

}  // namespace

struct hppb200_multi {
  std::vector<std::unique_ptr<Worker>> workers;
  std::mutex callMu;  // one sharded call at a time per handle
  ~hppb200_multi() {
    for (auto& w : workers) {
      if (w->thread.joinable()) {
        { std::lock_guard<std::mutex> lock(w->mu); w->quit = true; }
        w->cv.notify_all();
        w->thread.join();
      }
      if (w->ws) hppb200_workspace_destroy(w->ws);
    }
  }
  // runs fn(rank, begin, count) on every worker with a non-empty range; first failure wins
  int run(size_t n, const std::function<int(Worker&, size_t, size_t)>& fn) {
    std::lock_guard<std::mutex> lock(callMu);
    const size_t N = workers.size();
    std::vector<char> used(N, 0);
    for (size_t r = 0; r < N; ++r) {
      const size_t b = n * r / N, e = n * (r + 1) / N;
      if (e == b) continue;
      Worker* w = workers[r].get();
      used[r] = 1;
      w->submit([=]() { return fn(*w, b, e - b); });
    }
    int rc = HPPB200_OK;
    for (size_t r = 0; r < N; ++r) {
      if (!used[r]) continue;
      const int x = workers[r]->wait();
      if (x != HPPB200_OK && rc == HPPB200_OK) {
        rc = x;
        hppb200::setLastError("device " + std::to_string(workers[r]->device) + ": " + workers[r]->error);
      }
    }
    return rc;
  }
};

extern "C" {

int hppb200_multi_create(const hppb200_model* m, const int* devices, int ndevices, hppb200_multi** out) {
  if (!m || !out || ndevices < 0 || (ndevices > 0 && !devices)) { hppb200::setLastError("null argument"); return HPPB200_ERR_INVALID_ARGUMENT; }
  int count = hppb200_device_count();
  if (count == 0) { hppb200::setLastError("no CUDA device visible: the engine has no CPU fallback"); return HPPB200_ERR_CUDA; }
  std::unique_ptr<hppb200_multi> h(new hppb200_multi);
  const int N = ndevices > 0 ? ndevices : count;  // ndevices == 0: every visible device
  struct RestoreDevice {  // workspace creation makes its device current: leave the caller's as it was
    int prev = 0;
    RestoreDevice() { cudaGetDevice(&prev); }
    ~RestoreDevice() { cudaSetDevice(prev); }
  } restore;
  for (int r = 0; r < N; ++r) {
    std::unique_ptr<Worker> w(new Worker);
    w->device = ndevices > 0 ? devices[r] : r;
    const int rc = hppb200_workspace_create(m, w->device, &w->ws);
    if (rc != HPPB200_OK) return rc;  // (the partially built handle is destroyed with its workspaces)
    h->workers.push_back(std::move(w));
  }
  for (auto& w : h->workers) w->thread = std::thread([p = w.get()] { p->loop(); });
  *out = h.release();
  return HPPB200_OK;
}

void hppb200_multi_destroy(hppb200_multi* h) { delete h; }
int hppb200_multi_num_devices(const hppb200_multi* h) { return h ? (int)h->workers.size() : 0; }
hppb200_workspace* hppb200_multi_workspace(hppb200_multi* h, int rank) {
  return (h && rank >= 0 && rank < (int)h->workers.size()) ? h->workers[(size_t)rank]->ws : nullptr;
}

int hppb200_multi_validate_configs(hppb200_multi* h, const double* q, size_t n, uint32_t flags, uint8_t* valid, int32_t* first_pair,
                                   uint8_t* reason) {
  if (!h || (n && (!q || !valid))) { hppb200::setLastError("null argument"); return HPPB200_ERR_INVALID_ARGUMENT; }
  if (flags & HPPB200_DEVICE_POINTERS) { hppb200::setLastError("the multi-GPU entry points take host arrays"); return HPPB200_ERR_INVALID_ARGUMENT; }
  const size_t nq = (size_t)h->workers[0]->ws->impl.model->nq;
  return h->run(n, [=](Worker& w, size_t b, size_t c) {
    return hppb200_validate_configs(w.ws, q + b * nq, c, flags, valid + b, first_pair ? first_pair + b : nullptr, reason ? reason + b : nullptr);
  });
}

int hppb200_multi_validate_edges(hppb200_multi* h, int method, double param, int reverse, const double* q1, const double* q2, const double* L,
                                 size_t n, uint32_t flags, uint8_t* valid, double* t_fail, double* t_last_valid, int32_t* pair, uint8_t* reason) {
  if (!h || (n && (!q1 || !q2 || !L || !valid))) { hppb200::setLastError("null argument"); return HPPB200_ERR_INVALID_ARGUMENT; }
  if (flags & HPPB200_DEVICE_POINTERS) { hppb200::setLastError("the multi-GPU entry points take host arrays"); return HPPB200_ERR_INVALID_ARGUMENT; }
  const size_t nq = (size_t)h->workers[0]->ws->impl.model->nq;
  return h->run(n, [=](Worker& w, size_t b, size_t c) {
    return hppb200_validate_edges(w.ws, method, param, reverse, q1 + b * nq, q2 + b * nq, L + b, c, flags, valid + b, t_fail ? t_fail + b : nullptr,
                                  t_last_valid ? t_last_valid + b : nullptr, pair ? pair + b : nullptr, reason ? reason + b : nullptr);
  });
}

}  // extern "C"
