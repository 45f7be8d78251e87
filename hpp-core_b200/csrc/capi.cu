// C ABI of the engine (include/hppb200.h).  Status codes only; exceptions are caught here
// and turned into HPPB200_ERR_* + a thread-local message, mirroring the exception types
// the reference throws (std::invalid_argument / std::runtime_error).
// There is no CPU fallback: compute entry points require a CUDA device.
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "engine.hpp"
#include "knn_state.hpp"

namespace hppb200 {

static thread_local std::string g_last_error;
void setLastError(const std::string& s) { g_last_error = s; }

Workspace::Workspace(const Model* m, int dev) : model(m), device(dev) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
    throw std::runtime_error("no CUDA device visible: the engine has no CPU fallback");
  if (dev < 0 || dev >= count) throw std::invalid_argument("device index out of range");
  HPPB200_CUDA_CHECK(cudaSetDevice(dev));
  cudaDeviceProp prop;
  HPPB200_CUDA_CHECK(cudaGetDeviceProperties(&prop, dev));
  num_sms = prop.multiProcessorCount;
  HPPB200_CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  HPPB200_CUDA_CHECK(cudaEventCreate(&ev0));
  HPPB200_CUDA_CHECK(cudaEventCreate(&ev1));
  HPPB200_CUDA_CHECK(cudaMalloc((void**)&d_stats, sizeof(DeviceStats)));
  HPPB200_CUDA_CHECK(cudaMemset(d_stats, 0, sizeof(DeviceStats)));
  view = m->deviceView(dev);
  if (const char* v = std::getenv("HPPB200_ENGINE")) { use_engine = std::atoi(v) != 0; if (use_engine) engine_kind = std::atoi(v); }
  if (const char* v = std::getenv("HPPB200_WF_MIN")) wavefront_min = (size_t)std::atoll(v);
  if (const char* v = std::getenv("HPPB200_WF_TAIL")) wavefront_tail = std::atol(v);
  if (const char* v = std::getenv("HPPB200_TAIL")) use_tail = std::atoi(v) != 0;
  if (const char* v = std::getenv("HPPB200_TAIL_SMEM")) tail_smem = std::atoi(v) != 0;
  if (const char* v = std::getenv("HPPB200_TAIL_MAX")) tail_max = (size_t)std::atoll(v);
  if (const char* v = std::getenv("HPPB200_PROG_TAIL_MAX")) prog_tail_max = (size_t)std::atoll(v);
  if (const char* v = std::getenv("HPPB200_FIRST_SAMPLE_SCREEN")) first_sample_screen = std::atoi(v) != 0;
  if (const char* v = std::getenv("HPPB200_FS_MIN")) first_sample_min = (size_t)std::atoll(v);
  if (const char* v = std::getenv("HPPB200_FS_ORDER")) first_sample_order = std::atoi(v) != 0;
  if (const char* v = std::getenv("HPPB200_WF_SLOTS")) wavefront_slots = (size_t)std::atoll(v);
  if (const char* v = std::getenv("HPPB200_WF_LPT")) wavefront_lpt = std::atoi(v) != 0;
  if (const char* v = std::getenv("HPPB200_ENGINE_ALPHA")) engine_alpha = std::atof(v);
  if (const char* v = std::getenv("HPPB200_ENGINE_THREADS")) engine_threads = std::atoi(v);
}

Workspace::~Workspace() {
  cudaSetDevice(device);
  if (own_stream && stream) cudaStreamDestroy(stream);
  if (ev0) cudaEventDestroy(ev0);
  if (ev1) cudaEventDestroy(ev1);
  if (h_njobs) cudaFreeHost(h_njobs);
  if (h_small) cudaFreeHost(h_small);
  if (h_fs) cudaFreeHost(h_fs);
  if (edge_copy_stream) cudaStreamDestroy(edge_copy_stream);
  if (ev_edge_in) cudaEventDestroy(ev_edge_in);
  if (ev_edge_q1) cudaEventDestroy(ev_edge_q1);
  if (stream2) cudaStreamDestroy(stream2);
  if (stream3) cudaStreamDestroy(stream3);
  if (stream4) cudaStreamDestroy(stream4);
  if (stream5) cudaStreamDestroy(stream5);
  for (cudaEvent_t e : evSlice) cudaEventDestroy(e);
  if (evFork) cudaEventDestroy(evFork);
  if (evJoin) cudaEventDestroy(evJoin);
  if (d_stats) cudaFree(d_stats);
}

void Workspace::syncStats() {
  DeviceStats h;
  HPPB200_CUDA_CHECK(cudaMemcpyAsync(&h, d_stats, sizeof(h), cudaMemcpyDeviceToHost, stream));
  HPPB200_CUDA_CHECK(cudaStreamSynchronize(stream));
  stats.pair_tests = h.pair_tests;
  stats.gjk_iterations = h.gjk_iterations;
  stats.edge_samples = h.edge_samples;
  stats.iter_cap_hits = h.iter_cap_hits;  // (second-pass overflows are counted here by the kernel)
  stats.engine_ticks = h.engine_ticks;
  stats.discarded_tests = h.discarded_tests;
}

template <typename F>
static int guarded(F&& f) {
  try {
    f();
    return HPPB200_OK;
  } catch (const std::invalid_argument& e) {
    setLastError(e.what());
    return HPPB200_ERR_INVALID_ARGUMENT;
  } catch (const std::runtime_error& e) {
    setLastError(e.what());
    return std::strncmp(e.what(), "CUDA", 4) == 0 || std::strncmp(e.what(), "no CUDA", 7) == 0 ? HPPB200_ERR_CUDA
                                                                                            : HPPB200_ERR_RUNTIME;
  } catch (const std::exception& e) {
    setLastError(e.what());
    return HPPB200_ERR_RUNTIME;
  }
}

struct Timed {
  Workspace* ws;
  explicit Timed(Workspace* w) : ws(w) { cudaEventRecord(ws->ev0, ws->stream); }
  void stop() {
    cudaEventRecord(ws->ev1, ws->stream);
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, ws->ev0, ws->ev1);
    ws->stats.last_kernel_ms = ms;
  }
};

}  // namespace hppb200

using namespace hppb200;

extern "C" {

int hppb200_abi_version(void) { return HPPB200_ABI_VERSION; }
const char* hppb200_last_error(void) { return g_last_error.c_str(); }
int hppb200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

int hppb200_measure_fp64_peak(int device, double* tflops) {
  return guarded([&] {
    if (!tflops) throw std::invalid_argument("null argument");
    if (hppb200_device_count() == 0) throw std::runtime_error("no CUDA device visible: the engine has no CPU fallback");
    *tflops = measureFp64Peak(device);
  });
}

int hppb200_model_create(const hppb200_model_desc* desc, hppb200_model** out) {
  return guarded([&] {
    if (!desc || !out) throw std::invalid_argument("null argument");
    *out = new hppb200_model(*desc);
  });
}
void hppb200_model_destroy(hppb200_model* m) { delete m; }
int hppb200_model_nq(const hppb200_model* m) { return m->impl.nq; }
int hppb200_model_nv(const hppb200_model* m) { return m->impl.nv; }
int hppb200_model_num_config_pairs(const hppb200_model* m) { return (int)m->impl.cPairs.size(); }
int hppb200_model_get_config_pairs(const hppb200_model* m, int32_t* a, int32_t* b, double* margin) {
  for (size_t i = 0; i < m->impl.cPairs.size(); ++i) {
    if (a) a[i] = m->impl.cPairs[i].a;
    if (b) b[i] = m->impl.cPairs[i].b;
    if (margin) margin[i] = m->impl.cPairs[i].margin;
  }
  return HPPB200_OK;
}
int hppb200_model_num_parameterized_pairs(const hppb200_model* m) { return m->impl.nParameterized; }
int hppb200_model_num_body_pairs(const hppb200_model* m) { return (int)m->impl.bodyPairs.size(); }
int hppb200_model_get_body_pairs(const hppb200_model* m, int32_t* ja, int32_t* jb, int32_t* pair_begin,
                                 int32_t* coef_begin, int32_t* rcoef_begin) {
  int pb = 0, cb = 0, rb = 0;
  size_t i = 0;
  for (const BodyPairHost& b : m->impl.bodyPairs) {
    if (ja) ja[i] = b.joint_a;
    if (jb) jb[i] = b.joint_b;
    if (pair_begin) pair_begin[i] = pb;
    if (coef_begin) coef_begin[i] = cb;
    if (rcoef_begin) rcoef_begin[i] = rb;
    pb += (int)b.pairs.size();
    cb += (int)b.coefs.size();
    rb += (int)b.rcoefs.size();
    ++i;
  }
  if (pair_begin) pair_begin[i] = pb;
  if (coef_begin) coef_begin[i] = cb;
  if (rcoef_begin) rcoef_begin[i] = rb;
  return HPPB200_OK;
}
int hppb200_model_num_body_geom_pairs(const hppb200_model* m) {
  size_t n = 0;
  for (const BodyPairHost& b : m->impl.bodyPairs) n += b.pairs.size();
  return (int)n;
}
int hppb200_model_get_body_geom_pairs(const hppb200_model* m, int32_t* a, int32_t* b, double* margin) {
  size_t k = 0;
  for (const BodyPairHost& bp : m->impl.bodyPairs)
    for (const DevPair& p : bp.pairs) {
      if (a) a[k] = p.a;
      if (b) b[k] = p.b;
      if (margin) margin[k] = p.margin;
      ++k;
    }
  return HPPB200_OK;
}
int hppb200_model_num_coefficients(const hppb200_model* m, int reverse) {
  size_t n = 0;
  for (const BodyPairHost& b : m->impl.bodyPairs) n += reverse ? b.rcoefs.size() : b.coefs.size();
  return (int)n;
}
int hppb200_model_get_coefficients(const hppb200_model* m, int reverse, int32_t* joint, double* value) {
  size_t k = 0;
  for (const BodyPairHost& b : m->impl.bodyPairs)
    for (auto& c : (reverse ? b.rcoefs : b.coefs)) {
      if (joint) joint[k] = c.first;
      if (value) value[k] = c.second;
      ++k;
    }
  return HPPB200_OK;
}
int hppb200_model_get_body_radii(const hppb200_model* m, double* radius) {
  for (size_t i = 0; i < m->impl.radii.size(); ++i) radius[i] = m->impl.radii[i];
  return HPPB200_OK;
}

int hppb200_workspace_create(const hppb200_model* m, int device, hppb200_workspace** out) {
  return guarded([&] {
    if (!m || !out) throw std::invalid_argument("null argument");
    *out = new hppb200_workspace(&m->impl, device);
    try { noallocSetup(device); } catch (...) {}  // set-up of hppb200_validate_configs_noalloc (it reports a failure itself)
  });
}
void hppb200_workspace_destroy(hppb200_workspace* ws) { delete ws; }
int hppb200_get_stats(const hppb200_workspace* ws, hppb200_stats* out) {
  return guarded([&] {
    const_cast<hppb200_workspace*>(ws)->impl.syncStats();
    *out = ws->impl.stats;
  });
}
int hppb200_workspace_set_stream(hppb200_workspace* ws, void* cuda_stream) {
  return guarded([&] {
    Workspace& w = ws->impl;
    if (w.own_stream && w.stream) cudaStreamDestroy(w.stream);
    w.stream = (cudaStream_t)cuda_stream;
    w.own_stream = false;
  });
}
int hppb200_workspace_set_max_iterations(hppb200_workspace* ws, uint32_t max_iter) {
  ws->impl.max_iterations = max_iter ? max_iter : 1u;
  return HPPB200_OK;
}

int hppb200_validate_configs(hppb200_workspace* wsp, const double* q, size_t n, uint32_t flags, uint8_t* valid,
                             int32_t* first_pair, uint8_t* reason) {
  return guarded([&] {
    if (!wsp || (n && (!q || !valid))) throw std::invalid_argument("null argument");
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq;
    if (flags & HPPB200_DEVICE_POINTERS) {
      Timed tm(ws);
      launchValidateConfigs(ws, q, n, flags, valid, first_pair, reason);
      tm.stop();  // (synchronises)
      finishValidateConfigs(ws, q, n, flags, valid, first_pair, reason);
      return;
    }
    if (n == 0) return;
    static const size_t smallMax = [] { const char* v = std::getenv("HPPB200_SMALL_CALL_MAX"); return v ? (size_t)std::atoll(v) : (size_t)32; }();
    if (n <= smallMax && n * nq * sizeof(double) + 16 * n + 64 <= SMALL_STAGE_BYTES / 2) {
      // A planner's one-item call (ConfigValidation::validate(q)): one pinned block up, the CTA-cooperative kernel (pairs spread
      // over the warps: a lone configuration does not wait for one thread to walk all of them), one block down, one synchronisation.
      if (!ws->h_small) HPPB200_CUDA_CHECK(cudaHostAlloc((void**)&ws->h_small, SMALL_STAGE_BYTES, cudaHostAllocDefault));
      ws->d_small.reserve(SMALL_STAGE_BYTES);
      const size_t inBytes = n * nq * sizeof(double);
      const size_t oFirst = 0, oValid = n * sizeof(int32_t), oReason = oValid + n, outBytes = oReason + n;
      const size_t outOff = SMALL_STAGE_BYTES / 2;
      std::memcpy(ws->h_small, q, inBytes);
      Timed tm(ws);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_small.p, ws->h_small, inBytes, cudaMemcpyHostToDevice, ws->stream));
      unsigned char* dout = ws->d_small.p + outOff;
      launchValidateConfigsSmall(ws, (const double*)ws->d_small.p, n, flags, dout + oValid, (int32_t*)(dout + oFirst), dout + oReason);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->h_small + outOff, dout, outBytes, cudaMemcpyDeviceToHost, ws->stream));
      tm.stop();
      std::memcpy(valid, ws->h_small + outOff + oValid, n);
      if (first_pair) std::memcpy(first_pair, ws->h_small + outOff + oFirst, n * sizeof(int32_t));
      if (reason) std::memcpy(reason, ws->h_small + outOff + oReason, n);
      return;
    }
    // slices on two streams: copies overlap the kernels (kernels_config.cu, launchValidateConfigsHost)
    HPPB200_CUDA_CHECK(cudaEventRecord(ws->ev0, ws->stream));
    launchValidateConfigsHost(ws, q, n, flags, valid, first_pair, reason);  // records ev1, returns synchronised
    float ms = 0;
    cudaEventElapsedTime(&ms, ws->ev0, ws->ev1);
    ws->stats.last_kernel_ms = ms;
  });
}

int hppb200_validate_configs_noalloc(hppb200_workspace* wsp, const double* q, size_t n, uint32_t flags, uint8_t* valid,
                                     int32_t* first_pair, uint8_t* reason) {
  return guarded([&] {
    if (!wsp) throw std::invalid_argument("null argument");
    if (!(flags & HPPB200_DEVICE_POINTERS))
      throw std::invalid_argument("hppb200_validate_configs_noalloc takes device pointers only (HPPB200_DEVICE_POINTERS): staging host buffers would allocate");
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    Timed tm(ws);
    launchValidateConfigsNoAlloc(ws, q, n, flags, valid, first_pair, reason);
    tm.stop();
  });
}

int hppb200_geometry_placements(hppb200_workspace* wsp, const double* q, size_t n, uint32_t flags, double* oMg) {
  return guarded([&] {
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq, per = (size_t)ws->model->geoms.size() * 12;
    if (flags & HPPB200_DEVICE_POINTERS) { launchGeometryPlacements(ws, q, n, oMg); HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream)); return; }
    ws->d_q.reserve(n * nq);
    ws->d_scratch.reserve(n * per);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, q, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    launchGeometryPlacements(ws, ws->d_q.p, n, ws->d_scratch.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(oMg, ws->d_scratch.p, n * per * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  });
}

int hppb200_pair_distances(hppb200_workspace* wsp, const double* q, size_t n, uint32_t flags, double* dist) {
  return guarded([&] {
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq, per = ws->model->cPairs.size();
    if (flags & HPPB200_DEVICE_POINTERS) { launchPairDistances(ws, q, n, dist); HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream)); return; }
    ws->d_q.reserve(n * nq);
    ws->d_scratch.reserve(n * per);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, q, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    launchPairDistances(ws, ws->d_q.p, n, ws->d_scratch.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(dist, ws->d_scratch.p, n * per * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  });
}

int hppb200_all_collisions(hppb200_workspace* wsp, const double* q, size_t n, uint32_t flags, uint8_t* valid, int32_t* first_pair,
                           uint32_t* mask, int32_t* count) {
  return guarded([&] {
    if (!wsp || (n && (!q || !valid || !mask))) throw std::invalid_argument("null argument");
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq, words = (ws->model->cPairs.size() + 31) / 32;
    if (flags & HPPB200_DEVICE_POINTERS) {
      launchAllCollisions(ws, q, n, valid, first_pair, mask, count);
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
      return;
    }
    ws->d_q.reserve(n * nq); ws->d_valid.reserve(n); ws->d_pair.reserve(n); ws->d_mask.reserve(n * words); ws->d_ncoll.reserve(n);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, q, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    launchAllCollisions(ws, ws->d_q.p, n, ws->d_valid.p, ws->d_pair.p, ws->d_mask.p, ws->d_ncoll.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(valid, ws->d_valid.p, n, cudaMemcpyDeviceToHost, ws->stream));
    if (first_pair) HPPB200_CUDA_CHECK(cudaMemcpyAsync(first_pair, ws->d_pair.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(mask, ws->d_mask.p, n * words * sizeof(uint32_t), cudaMemcpyDeviceToHost, ws->stream));
    if (count) HPPB200_CUDA_CHECK(cudaMemcpyAsync(count, ws->d_ncoll.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  });
}

int hppb200_model_num_distance_pairs(const hppb200_model* m) { return (int)m->impl.dboPairs.size(); }
int hppb200_distance_between_objects(hppb200_workspace* wsp, const double* q, size_t n, uint32_t flags, double* dist) {
  return guarded([&] {
    if (!wsp || (n && (!q || !dist))) throw std::invalid_argument("null argument");
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq, per = ws->model->dboPairs.size();
    if (flags & HPPB200_DEVICE_POINTERS) {
      launchDistanceBetweenObjects(ws, q, n, dist);
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
      return;
    }
    ws->d_q.reserve(n * nq); ws->d_tfail.reserve(n * per + 1);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, q, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    launchDistanceBetweenObjects(ws, ws->d_q.p, n, ws->d_tfail.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(dist, ws->d_tfail.p, n * per * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  });
}

int hppb200_straight_lengths(hppb200_workspace* wsp, const double* q1, const double* q2, size_t n, uint32_t flags,
                             double* L) {
  return guarded([&] {
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq;
    if (flags & HPPB200_DEVICE_POINTERS) { launchStraightLengths(ws, q1, q2, n, L); HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream)); return; }
    if (n && n <= 32 && (2 * nq + 1) * n * sizeof(double) <= SMALL_STAGE_BYTES / 2) {  // one-item call: one block up, one down (see hppb200_validate_edges)
      if (!ws->h_small) HPPB200_CUDA_CHECK(cudaHostAlloc((void**)&ws->h_small, SMALL_STAGE_BYTES, cudaHostAllocDefault));
      ws->d_small.reserve(SMALL_STAGE_BYTES);
      const size_t qBytes = n * nq * sizeof(double), outOff = SMALL_STAGE_BYTES / 2;
      std::memcpy(ws->h_small, q1, qBytes);
      std::memcpy(ws->h_small + qBytes, q2, qBytes);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_small.p, ws->h_small, 2 * qBytes, cudaMemcpyHostToDevice, ws->stream));
      launchStraightLengths(ws, (const double*)ws->d_small.p, (const double*)(ws->d_small.p + qBytes), n, (double*)(ws->d_small.p + outOff));
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->h_small + outOff, ws->d_small.p + outOff, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
      HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
      std::memcpy(L, ws->h_small + outOff, n * sizeof(double));
      return;
    }
    ws->d_q.reserve(n * nq);
    ws->d_q2.reserve(n * nq);
    ws->d_L.reserve(n);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, q1, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q2.p, q2, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    launchStraightLengths(ws, ws->d_q.p, ws->d_q2.p, n, ws->d_L.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(L, ws->d_L.p, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  });
}

int hppb200_path_eval(hppb200_workspace* wsp, const double* q1, const double* q2, const double* L, const double* t, size_t n,
                      uint32_t flags, double* q) {
  return guarded([&] {
    if (!wsp || (n && (!q1 || !q2 || !L || !t || !q))) throw std::invalid_argument("null argument");
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq;
    if (flags & HPPB200_DEVICE_POINTERS) { launchPathEval(ws, q1, q2, L, t, n, q); HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream)); return; }
    ws->d_q.reserve(n * nq);
    ws->d_q2.reserve(n * nq);
    ws->d_L.reserve(n);
    ws->d_tfail.reserve(n);
    ws->d_scratch.reserve(n * nq);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, q1, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q2.p, q2, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_L.p, L, n * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_tfail.p, t, n * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    launchPathEval(ws, ws->d_q.p, ws->d_q2.p, ws->d_L.p, ws->d_tfail.p, n, ws->d_scratch.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(q, ws->d_scratch.p, n * nq * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
  });
}

int hppb200_validate_edges(hppb200_workspace* wsp, int method, double param, int reverse, const double* q1,
                           const double* q2, const double* L, size_t n, uint32_t flags, uint8_t* valid, double* t_fail,
                           double* t_last_valid, int32_t* pair, uint8_t* reason) {
  return guarded([&] {
    if (!wsp || (n && (!q1 || !q2 || !L || !valid))) throw std::invalid_argument("null argument");
    // constructor checks of the reference validators
    if (method == HPPB200_METHOD_PROGRESSIVE && !(param > 0))
      throw std::runtime_error("tolerance should be positive for progressive continuous validation.");  // progressive.cc:138-145
    if (method == HPPB200_METHOD_DICHOTOMY && param != 0)
      throw std::runtime_error("Dichotomy path validation method does not support penetration.");  // dichotomy.cc:139-148
    if (method == HPPB200_METHOD_DISCRETIZED && !(param > 0)) throw std::invalid_argument("step size must be positive");
    if (method < HPPB200_METHOD_DISCRETIZED || method > HPPB200_METHOD_DICHOTOMY) throw std::invalid_argument("unknown method");
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq;
    if (flags & HPPB200_DEVICE_POINTERS) {
      Timed tm(ws);
      launchValidateEdges(ws, method, param, reverse, q1, q2, L, n, flags, valid, t_fail, t_last_valid, pair, reason);
      tm.stop();
      return;
    }
    static const size_t smallMax = [] { const char* v = std::getenv("HPPB200_SMALL_CALL_MAX"); return v ? (size_t)std::atoll(v) : (size_t)32; }();
    if (n && n <= smallMax && (2 * nq + 1) * n * sizeof(double) + 32 * n + 64 <= SMALL_STAGE_BYTES / 2) {
      // A planner's one-item call (PathValidation::validate(path)): the two ends and the length in one pinned block and one H2D,
      // the five outputs in one device block and one D2H.
      if (!ws->h_small) HPPB200_CUDA_CHECK(cudaHostAlloc((void**)&ws->h_small, SMALL_STAGE_BYTES, cudaHostAllocDefault));
      ws->d_small.reserve(SMALL_STAGE_BYTES);
      const size_t qBytes = n * nq * sizeof(double), inBytes = 2 * qBytes + n * sizeof(double);
      const size_t oTf = 0, oTl = n * sizeof(double), oPair = 2 * n * sizeof(double), oValid = oPair + n * sizeof(int32_t), oReason = oValid + n,
                   outBytes = oReason + n, outOff = SMALL_STAGE_BYTES / 2;
      std::memcpy(ws->h_small, q1, qBytes);
      std::memcpy(ws->h_small + qBytes, q2, qBytes);
      std::memcpy(ws->h_small + 2 * qBytes, L, n * sizeof(double));
      Timed tm(ws);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_small.p, ws->h_small, inBytes, cudaMemcpyHostToDevice, ws->stream));
      unsigned char* din = ws->d_small.p;
      unsigned char* dout = ws->d_small.p + outOff;
      launchValidateEdges(ws, method, param, reverse, (const double*)din, (const double*)(din + qBytes), (const double*)(din + 2 * qBytes), n, flags,
                          dout + oValid, (double*)(dout + oTf), (double*)(dout + oTl), (int32_t*)(dout + oPair), dout + oReason);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->h_small + outOff, dout, outBytes, cudaMemcpyDeviceToHost, ws->stream));
      tm.stop();
      const unsigned char* ho = ws->h_small + outOff;
      std::memcpy(valid, ho + oValid, n);
      if (t_fail) std::memcpy(t_fail, ho + oTf, n * sizeof(double));
      if (t_last_valid) std::memcpy(t_last_valid, ho + oTl, n * sizeof(double));
      if (pair) std::memcpy(pair, ho + oPair, n * sizeof(int32_t));
      if (reason) std::memcpy(reason, ho + oReason, n);
      return;
    }
    ws->d_q.reserve(n * nq);
    ws->d_q2.reserve(n * nq);
    ws->d_L.reserve(n);
    ws->d_valid.reserve(n);
    ws->d_reason.reserve(n);
    ws->d_pair.reserve(n);
    ws->d_tfail.reserve(n);
    ws->d_tlast.reserve(n);
    Timed tm(ws);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, q1, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    // A batch the first-sample screen will look at (forward: it needs q1 only): q2 and L on their own stream, under the screen
    // (the 30-joint model of BASELINE configs[3]: 240 MB of the 511 MB a 1 M-edge call brings in)
    const bool split = ws->first_sample_screen && method != HPPB200_METHOD_DISCRETIZED && !reverse && n >= ws->first_sample_min;
    cudaStream_t s2 = ws->stream;
    if (split) {
      if (!ws->edge_copy_stream) {
        HPPB200_CUDA_CHECK(cudaStreamCreateWithFlags(&ws->edge_copy_stream, cudaStreamNonBlocking));
        HPPB200_CUDA_CHECK(cudaEventCreateWithFlags(&ws->ev_edge_in, cudaEventDisableTiming));
        HPPB200_CUDA_CHECK(cudaEventCreateWithFlags(&ws->ev_edge_q1, cudaEventDisableTiming));
      }
      s2 = ws->edge_copy_stream;  // (nothing of an earlier call is in flight: every call returns synchronised)
      // behind q1, not beside it: two H2D copies in flight share the link and q1 would arrive with q2
      HPPB200_CUDA_CHECK(cudaEventRecord(ws->ev_edge_q1, ws->stream));
      HPPB200_CUDA_CHECK(cudaStreamWaitEvent(s2, ws->ev_edge_q1, 0));
    }
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q2.p, q2, n * nq * sizeof(double), cudaMemcpyHostToDevice, s2));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_L.p, L, n * sizeof(double), cudaMemcpyHostToDevice, s2));
    ws->edge_inputs_ready = nullptr;
    if (split) {
      HPPB200_CUDA_CHECK(cudaEventRecord(ws->ev_edge_in, s2));
      ws->edge_inputs_ready = ws->ev_edge_in;
    }
    try {
      launchValidateEdges(ws, method, param, reverse, ws->d_q.p, ws->d_q2.p, ws->d_L.p, n, flags, ws->d_valid.p,
                          ws->d_tfail.p, ws->d_tlast.p, ws->d_pair.p, ws->d_reason.p);
    } catch (...) {
      ws->edge_inputs_ready = nullptr;
      if (split) cudaStreamSynchronize(s2);
      throw;
    }
    if (ws->edge_inputs_ready) {  // (not consumed: cannot happen, kept for safety)
      HPPB200_CUDA_CHECK(cudaStreamWaitEvent(ws->stream, ws->edge_inputs_ready, 0));
      ws->edge_inputs_ready = nullptr;
    }
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(valid, ws->d_valid.p, n, cudaMemcpyDeviceToHost, ws->stream));
    if (t_fail) HPPB200_CUDA_CHECK(cudaMemcpyAsync(t_fail, ws->d_tfail.p, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    if (t_last_valid) HPPB200_CUDA_CHECK(cudaMemcpyAsync(t_last_valid, ws->d_tlast.p, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    if (pair) HPPB200_CUDA_CHECK(cudaMemcpyAsync(pair, ws->d_pair.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    if (reason) HPPB200_CUDA_CHECK(cudaMemcpyAsync(reason, ws->d_reason.p, n, cudaMemcpyDeviceToHost, ws->stream));
    tm.stop();
  });
}

int hppb200_validate_spline_edges(hppb200_workspace* wsp, int method, double param, int reverse, const double* base,
                                  const double* params, const double* L, size_t n, uint32_t flags, uint8_t* valid, double* t_fail,
                                  double* t_last_valid, int32_t* pair, uint8_t* reason) {
  return guarded([&] {
    if (!wsp || (n && (!base || !params || !L || !valid))) throw std::invalid_argument("null argument");
    if (method == HPPB200_METHOD_PROGRESSIVE && !(param > 0))
      throw std::runtime_error("tolerance should be positive for progressive continuous validation.");  // progressive.cc:138-145
    if (method == HPPB200_METHOD_DICHOTOMY && param != 0)
      throw std::runtime_error("Dichotomy path validation method does not support penetration.");  // dichotomy.cc:139-148
    if (method == HPPB200_METHOD_DISCRETIZED && !(param > 0)) throw std::invalid_argument("step size must be positive");
    if (method < HPPB200_METHOD_DISCRETIZED || method > HPPB200_METHOD_DICHOTOMY) throw std::invalid_argument("unknown method");
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq;
    if (ws->model->nq != ws->model->nv)
      throw std::invalid_argument("spline paths: vector-space configuration spaces only (no free-flyer joints)");
    if (flags & HPPB200_DEVICE_POINTERS) {
      Timed tm(ws);
      launchValidateSplineEdges(ws, method, param, reverse, base, params, L, n, flags, valid, t_fail, t_last_valid, pair, reason);
      tm.stop();
      return;
    }
    ws->d_q.reserve(n * nq);
    ws->d_q2.reserve(n * 4 * nq);
    ws->d_L.reserve(n);
    ws->d_valid.reserve(n);
    ws->d_reason.reserve(n);
    ws->d_pair.reserve(n);
    ws->d_tfail.reserve(n);
    ws->d_tlast.reserve(n);
    Timed tm(ws);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, base, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q2.p, params, n * 4 * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_L.p, L, n * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
    launchValidateSplineEdges(ws, method, param, reverse, ws->d_q.p, ws->d_q2.p, ws->d_L.p, n, flags, ws->d_valid.p, ws->d_tfail.p,
                              ws->d_tlast.p, ws->d_pair.p, ws->d_reason.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(valid, ws->d_valid.p, n, cudaMemcpyDeviceToHost, ws->stream));
    if (t_fail) HPPB200_CUDA_CHECK(cudaMemcpyAsync(t_fail, ws->d_tfail.p, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    if (t_last_valid) HPPB200_CUDA_CHECK(cudaMemcpyAsync(t_last_valid, ws->d_tlast.p, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    if (pair) HPPB200_CUDA_CHECK(cudaMemcpyAsync(pair, ws->d_pair.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    if (reason) HPPB200_CUDA_CHECK(cudaMemcpyAsync(reason, ws->d_reason.p, n, cudaMemcpyDeviceToHost, ws->stream));
    tm.stop();
  });
}

int hppb200_validate_edge_chains(hppb200_workspace* wsp, int method, double param, int reverse, const double* q1, const double* q2,
                                 const double* L, size_t n_edges, const int32_t* chain_begin, size_t n_chains, uint32_t flags,
                                 uint8_t* valid, double* t_fail, double* t_last_valid, int32_t* pair, uint8_t* reason,
                                 int32_t* first_bad_edge) {
  return guarded([&] {
    if (!wsp || (n_chains && (!chain_begin || !valid)) || (n_edges && (!q1 || !q2 || !L))) throw std::invalid_argument("null argument");
    if (method == HPPB200_METHOD_DISCRETIZED)  // Discretized walks a PathVector as ONE path (discretized.cc:55-107): no recursion to replay
      throw std::runtime_error("edge chains replay the PathVector recursion of the continuous validations only");
    if (method == HPPB200_METHOD_PROGRESSIVE && !(param > 0))
      throw std::runtime_error("tolerance should be positive for progressive continuous validation.");
    if (method == HPPB200_METHOD_DICHOTOMY && param != 0) throw std::runtime_error("Dichotomy path validation method does not support penetration.");
    if (method != HPPB200_METHOD_PROGRESSIVE && method != HPPB200_METHOD_DICHOTOMY) throw std::invalid_argument("unknown method");
    if (n_edges > 0x7fffffffull) throw std::invalid_argument("at most 2^31-1 edges per call");
    Workspace* ws = &wsp->impl;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    const size_t nq = (size_t)ws->model->nq, n = n_edges;
    const bool dev = (flags & HPPB200_DEVICE_POINTERS) != 0;
    ws->d_valid.reserve(n); ws->d_reason.reserve(n); ws->d_pair.reserve(n); ws->d_tfail.reserve(n); ws->d_tlast.reserve(n);
    Timed tm(ws);
    const double *dq1 = q1, *dq2 = q2, *dL = L;
    const int32_t* dcb = chain_begin;
    if (!dev) {
      ws->d_q.reserve(n * nq); ws->d_q2.reserve(n * nq); ws->d_L.reserve(n); ws->d_chain.reserve(n_chains + 1);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q.p, q1, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_q2.p, q2, n * nq * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_L.p, L, n * sizeof(double), cudaMemcpyHostToDevice, ws->stream));
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_chain.p, chain_begin, (n_chains + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, ws->stream));
      dq1 = ws->d_q.p; dq2 = ws->d_q2.p; dL = ws->d_L.p; dcb = ws->d_chain.p;
    }
    if (n) launchValidateEdges(ws, method, param, reverse, dq1, dq2, dL, n, flags, ws->d_valid.p, ws->d_tfail.p, ws->d_tlast.p, ws->d_pair.p,
                               ws->d_reason.p);
    if (dev) {
      launchChainReduce(ws, dcb, n_chains, reverse, dL, ws->d_valid.p, ws->d_tfail.p, ws->d_tlast.p, ws->d_pair.p, ws->d_reason.p, valid, t_fail,
                        t_last_valid, pair, reason, first_bad_edge);
    } else {
      ws->d_cvalid.reserve(n_chains); ws->d_creason.reserve(n_chains); ws->d_cpair.reserve(n_chains); ws->d_cbad.reserve(n_chains);
      ws->d_ctfail.reserve(n_chains); ws->d_ctlast.reserve(n_chains);
      launchChainReduce(ws, dcb, n_chains, reverse, dL, ws->d_valid.p, ws->d_tfail.p, ws->d_tlast.p, ws->d_pair.p, ws->d_reason.p, ws->d_cvalid.p,
                        ws->d_ctfail.p, ws->d_ctlast.p, ws->d_cpair.p, ws->d_creason.p, ws->d_cbad.p);
      HPPB200_CUDA_CHECK(cudaMemcpyAsync(valid, ws->d_cvalid.p, n_chains, cudaMemcpyDeviceToHost, ws->stream));
      if (t_fail) HPPB200_CUDA_CHECK(cudaMemcpyAsync(t_fail, ws->d_ctfail.p, n_chains * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
      if (t_last_valid) HPPB200_CUDA_CHECK(cudaMemcpyAsync(t_last_valid, ws->d_ctlast.p, n_chains * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
      if (pair) HPPB200_CUDA_CHECK(cudaMemcpyAsync(pair, ws->d_cpair.p, n_chains * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
      if (reason) HPPB200_CUDA_CHECK(cudaMemcpyAsync(reason, ws->d_creason.p, n_chains, cudaMemcpyDeviceToHost, ws->stream));
      if (first_bad_edge) HPPB200_CUDA_CHECK(cudaMemcpyAsync(first_bad_edge, ws->d_cbad.p, n_chains * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    }
    tm.stop();
  });
}

int hppb200_validate_edges_indexed(hppb200_workspace* wsp, hppb200_knn* nn, int method, double param, int reverse, const int64_t* node_a,
                                   const int64_t* node_b, size_t n, uint32_t flags, uint8_t* valid, double* t_fail, double* t_last_valid,
                                   int32_t* pair, uint8_t* reason, double* length) {
  return guarded([&] {
    if (!wsp || !nn || (n && (!node_a || !node_b || !valid))) throw std::invalid_argument("null argument");
    if (flags & HPPB200_DEVICE_POINTERS) throw std::invalid_argument("hppb200_validate_edges_indexed takes host arrays");
    if (method == HPPB200_METHOD_PROGRESSIVE && !(param > 0))
      throw std::runtime_error("tolerance should be positive for progressive continuous validation.");
    if (method == HPPB200_METHOD_DICHOTOMY && param != 0) throw std::runtime_error("Dichotomy path validation method does not support penetration.");
    if (method == HPPB200_METHOD_DISCRETIZED && !(param > 0)) throw std::invalid_argument("step size must be positive");
    if (method < HPPB200_METHOD_DISCRETIZED || method > HPPB200_METHOD_DICHOTOMY) throw std::invalid_argument("unknown method");
    Workspace* ws = &wsp->impl;
    if (nn->device != ws->device || nn->model->nq != ws->model->nq) throw std::invalid_argument("the roadmap mirror and the workspace must share device and robot");
    if (n == 0) return;
    HPPB200_CUDA_CHECK(cudaSetDevice(ws->device));
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(nn->stream));  // appended nodes are in place
    const size_t nq = (size_t)ws->model->nq;
    ws->d_ia.reserve(n); ws->d_ib.reserve(n); ws->d_q.reserve(n * nq); ws->d_q2.reserve(n * nq); ws->d_L.reserve(n);
    ws->d_valid.reserve(n); ws->d_reason.reserve(n); ws->d_pair.reserve(n); ws->d_tfail.reserve(n); ws->d_tlast.reserve(n);
    ws->d_count.reserve(2);
    Timed tm(ws);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_ia.p, node_a, n * sizeof(int64_t), cudaMemcpyHostToDevice, ws->stream));
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(ws->d_ib.p, node_b, n * sizeof(int64_t), cudaMemcpyHostToDevice, ws->stream));
    unsigned* d_bad = (unsigned*)ws->d_count.p;
    HPPB200_CUDA_CHECK(cudaMemsetAsync(d_bad, 0, sizeof(unsigned), ws->stream));
    launchGatherEdgeEnds(ws->stream, nn->d_nodes, nn->n, (int)nq, (const int64_t*)ws->d_ia.p, (const int64_t*)ws->d_ib.p, n, ws->d_q.p, ws->d_q2.p, d_bad);
    unsigned h_bad = 0;
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(&h_bad, d_bad, sizeof(unsigned), cudaMemcpyDeviceToHost, ws->stream));
    launchStraightLengths(ws, ws->d_q.p, ws->d_q2.p, n, ws->d_L.p);  // steeringMethod::Straight: L = WeighedDistance(q1, q2)
    HPPB200_CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    if (h_bad) throw std::invalid_argument("node index out of range of the roadmap mirror");
    ws->stats.kernel_launches += 1;
    launchValidateEdges(ws, method, param, reverse, ws->d_q.p, ws->d_q2.p, ws->d_L.p, n, flags, ws->d_valid.p, ws->d_tfail.p, ws->d_tlast.p,
                        ws->d_pair.p, ws->d_reason.p);
    HPPB200_CUDA_CHECK(cudaMemcpyAsync(valid, ws->d_valid.p, n, cudaMemcpyDeviceToHost, ws->stream));
    if (t_fail) HPPB200_CUDA_CHECK(cudaMemcpyAsync(t_fail, ws->d_tfail.p, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    if (t_last_valid) HPPB200_CUDA_CHECK(cudaMemcpyAsync(t_last_valid, ws->d_tlast.p, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    if (pair) HPPB200_CUDA_CHECK(cudaMemcpyAsync(pair, ws->d_pair.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    if (reason) HPPB200_CUDA_CHECK(cudaMemcpyAsync(reason, ws->d_reason.p, n, cudaMemcpyDeviceToHost, ws->stream));
    if (length) HPPB200_CUDA_CHECK(cudaMemcpyAsync(length, ws->d_L.p, n * sizeof(double), cudaMemcpyDeviceToHost, ws->stream));
    tm.stop();
  });
}

// The helpers below launch on a caller-provided stream: the launch needs the stream's device current, whatever the calling
// thread's current device is (another library call may have changed it): take it from the data pointer, restore after.
namespace {
struct DeviceOfPointer {
  int prev = -1;
  explicit DeviceOfPointer(const void* p) {
    cudaPointerAttributes a{};
    HPPB200_CUDA_CHECK(cudaGetDevice(&prev));
    HPPB200_CUDA_CHECK(cudaPointerGetAttributes(&a, p));
    if (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged) throw std::invalid_argument("device pointers expected");
    if (a.device != prev) HPPB200_CUDA_CHECK(cudaSetDevice(a.device)); else prev = -1;
  }
  ~DeviceOfPointer() { if (prev >= 0) cudaSetDevice(prev); }
};
}  // namespace

// ---- exchange helpers of the sharded form (device pointers, asynchronous on `cuda_stream`) ----
int hppb200_pack_mask(void* cuda_stream, const uint8_t* valid, size_t n, uint32_t* bits) {
  return guarded([&] {
    if (n && (!valid || !bits)) throw std::invalid_argument("null argument");
    if (!n) return;
    DeviceOfPointer guard(valid);
    launchPackMask((cudaStream_t)cuda_stream, valid, n, bits);
  });
}
int hppb200_unpack_mask(void* cuda_stream, const uint32_t* bits, size_t n, uint8_t* valid) {
  return guarded([&] {
    if (n && (!valid || !bits)) throw std::invalid_argument("null argument");
    if (!n) return;
    DeviceOfPointer guard(valid);
    launchUnpackMask((cudaStream_t)cuda_stream, bits, n, valid);
  });
}
int hppb200_knn_merge_shards(void* cuda_stream, const double* dist, const int64_t* idx, int nshards, size_t nqueries, int k,
                             int64_t* out_idx, double* out_dist) {
  return guarded([&] {
    if (nqueries && (!dist || !idx || !out_idx || !out_dist)) throw std::invalid_argument("null argument");
    if (!nqueries) return;
    DeviceOfPointer guard(dist);
    launchKnnMergeShards((cudaStream_t)cuda_stream, dist, idx, nshards, nqueries, k, out_idx, out_dist);
  });
}

}  // extern "C"
