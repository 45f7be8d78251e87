// Fused exchange step of the sharded form: the validity mask of a rank's shard is bit-packed
// and written straight into EVERY rank's gather buffer over NVLink peer memory by one kernel, completion signalled by
// per-rank epoch flags in the receiver's memory; a second, one-warp kernel waits until all ranks' flags of this epoch
// have arrived.  It replaces hppb200_pack_mask + ncclAllGather (128 KiB per rank: a latency-bound collective, ~55 us
// per step at 2 GPUs against 0.81 ms of kernels) for processes on one node.
//
// Memory: one cudaMalloc block per rank, exported with cudaIpcGetMemHandle and opened by the peers:
//   [0, 1024)                 flags: uint32 epoch[world] (written by the peers, system scope)
//   [1024, 1024 + 2 * slot)   two gather slots (epoch parity), each world * words uint32, rank r's words at r * words
// Two slots: a rank can be at most one exchange ahead of the slowest (an exchange completes only when every rank has
// delivered), so the slot of epoch e is rewritten at epoch e + 2, after every rank has finished reading epoch e.
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "engine.hpp"

namespace hppb200 {
void setLastError(const std::string& s);
}

struct hppb200_exchange {
  int device = 0, rank = 0, world = 1;
  size_t max_words = 0;        // per rank and slot
  unsigned char* local = nullptr;
  std::vector<unsigned char*> peer;  // [world], peer[rank] == local
  std::vector<bool> opened;
  unsigned char** d_peer = nullptr;  // device copy of `peer`
  unsigned* d_ticket = nullptr;
  uint32_t epoch = 0;
};

namespace {
constexpr size_t EX_FLAGS_BYTES = 1024;  // epoch[64], then the time-out word
constexpr int EX_MAX_WORLD = 64;

template <class F>
int exGuard(F&& f) {
  try {
    f();
    return HPPB200_OK;
  } catch (const std::invalid_argument& e) {
    hppb200::setLastError(e.what());
    return HPPB200_ERR_INVALID_ARGUMENT;
  } catch (const std::exception& e) {
    hppb200::setLastError(e.what());
    return std::strncmp(e.what(), "CUDA", 4) == 0 ? HPPB200_ERR_CUDA : HPPB200_ERR_RUNTIME;
  }
}
}  // namespace

namespace hppb200 {

// valid [n] u8 -> this rank's packed words, stored into the slot of EVERY rank (peer memory); the last CTA to finish
// publishes the epoch in every rank's flag array.
__global__ void __launch_bounds__(256)
packScatterKernel(const uint8_t* __restrict__ valid, size_t n, size_t words, int rank, int world, unsigned char* const* __restrict__ peers,
                  size_t slotOff, uint32_t epoch, unsigned* ticket) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned b = __ballot_sync(0xffffffffu, i < n && valid[i] != 0);
  if ((threadIdx.x & 31) == 0 && i < n) {
    const size_t w = i >> 5;
    for (int r = 0; r < world; ++r) {
      uint32_t* slot = (uint32_t*)(peers[r] + slotOff);
      slot[(size_t)rank * words + w] = b;
    }
    __threadfence_system();  // this warp's words are visible to every GPU before the CTA takes its ticket
  }
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) {
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  if (threadIdx.x < world) {
    __threadfence_system();
    volatile uint32_t* flags = (volatile uint32_t*)peers[threadIdx.x];
    flags[rank] = epoch;  // (monotone: the waiter tests >=)
  }
  if (threadIdx.x == 0) *ticket = 0;
}

// One warp: lane r waits for rank r's epoch flag in THIS rank's memory (at most ~2 s: a peer that never delivers must not
// hang the stream; the time-out is recorded in the word behind the flags and reported by hppb200_exchange_status).
__global__ void waitFlagsKernel(unsigned char* local, int world, uint32_t epoch) {
  const int r = threadIdx.x;
  if (r < world) {
    const volatile uint32_t* flags = (const volatile uint32_t*)local;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    while (flags[r] < epoch) {
      __nanosleep(200);
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
      if (t1 - t0 > 2000000000ull) { ((volatile uint32_t*)local)[EX_MAX_WORLD] = 1u; break; }
    }
  }
  __threadfence_system();  // the words written before the flags are read after this point
}

}  // namespace hppb200

extern "C" {

int hppb200_exchange_create(int device, int rank, int world, size_t max_units, hppb200_exchange** out, unsigned char* handle64) {
  return exGuard([&] {
    if (!out || !handle64 || world < 1 || world > EX_MAX_WORLD || rank < 0 || rank >= world || max_units == 0)
      throw std::invalid_argument("hppb200_exchange_create: bad argument (1 <= world <= 64, 0 <= rank < world, max_units > 0)");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    HPPB200_CUDA_CHECK(cudaSetDevice(device));
    auto* ex = new hppb200_exchange;
    ex->device = device; ex->rank = rank; ex->world = world;
    ex->max_words = (max_units + 31) / 32;
    const size_t bytes = EX_FLAGS_BYTES + 2 * (size_t)world * ex->max_words * sizeof(uint32_t);
    try {
      HPPB200_CUDA_CHECK(cudaMalloc((void**)&ex->local, bytes));
      HPPB200_CUDA_CHECK(cudaMemset(ex->local, 0, bytes));
      HPPB200_CUDA_CHECK(cudaMalloc((void**)&ex->d_peer, world * sizeof(unsigned char*)));
      HPPB200_CUDA_CHECK(cudaMalloc((void**)&ex->d_ticket, sizeof(unsigned)));
      HPPB200_CUDA_CHECK(cudaMemset(ex->d_ticket, 0, sizeof(unsigned)));
      cudaIpcMemHandle_t h;
      HPPB200_CUDA_CHECK(cudaIpcGetMemHandle(&h, ex->local));
      std::memcpy(handle64, &h, 64);
      HPPB200_CUDA_CHECK(cudaDeviceSynchronize());
    } catch (...) {
      if (ex->local) cudaFree(ex->local);
      if (ex->d_peer) cudaFree(ex->d_peer);
      if (ex->d_ticket) cudaFree(ex->d_ticket);
      delete ex;
      throw;
    }
    ex->peer.assign(world, nullptr);
    ex->opened.assign(world, false);
    ex->peer[rank] = ex->local;
    *out = ex;
  });
}

int hppb200_exchange_open_peers(hppb200_exchange* ex, const unsigned char* handles /* [world][64], rank-major */) {
  return exGuard([&] {
    if (!ex || !handles) throw std::invalid_argument("null argument");
    HPPB200_CUDA_CHECK(cudaSetDevice(ex->device));
    for (int r = 0; r < ex->world; ++r) {
      if (r == ex->rank || ex->opened[r]) continue;
      cudaIpcMemHandle_t h;
      std::memcpy(&h, handles + 64 * (size_t)r, 64);
      void* p = nullptr;
      HPPB200_CUDA_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
      ex->peer[r] = (unsigned char*)p;
      ex->opened[r] = true;
    }
    HPPB200_CUDA_CHECK(cudaMemcpy(ex->d_peer, ex->peer.data(), ex->world * sizeof(unsigned char*), cudaMemcpyHostToDevice));
  });
}

int hppb200_exchange_mask(hppb200_exchange* ex, void* cuda_stream, const uint8_t* valid, size_t n, const uint32_t** gathered) {
  return exGuard([&] {
    if (!ex || !valid || !gathered) throw std::invalid_argument("null argument");
    const size_t words = (n + 31) / 32;
    if (n == 0 || words > ex->max_words) throw std::invalid_argument("hppb200_exchange_mask: n is 0 or exceeds the max_units the exchange was created with");
    for (int r = 0; r < ex->world; ++r)
      if (!ex->peer[r]) throw std::invalid_argument("hppb200_exchange_mask: hppb200_exchange_open_peers has not been called");
    HPPB200_CUDA_CHECK(cudaSetDevice(ex->device));
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    ex->epoch += 1;
    const size_t slotOff = EX_FLAGS_BYTES + (size_t)(ex->epoch & 1u) * ex->world * ex->max_words * sizeof(uint32_t);
    // (rank r's words at r * words of the slot: `words` of THIS call, the same n on every rank)
    hppb200::packScatterKernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(valid, n, words, ex->rank, ex->world, ex->d_peer, slotOff, ex->epoch,
                                                                                ex->d_ticket);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    hppb200::waitFlagsKernel<<<1, 64, 0, stream>>>(ex->local, ex->world, ex->epoch);
    HPPB200_CUDA_CHECK(cudaGetLastError());
    *gathered = (const uint32_t*)(ex->local + slotOff);
  });
}

int hppb200_exchange_status(hppb200_exchange* ex, int* timed_out) {
  return exGuard([&] {
    if (!ex || !timed_out) throw std::invalid_argument("null argument");
    HPPB200_CUDA_CHECK(cudaSetDevice(ex->device));
    HPPB200_CUDA_CHECK(cudaDeviceSynchronize());
    uint32_t w = 0;
    HPPB200_CUDA_CHECK(cudaMemcpy(&w, ex->local + EX_MAX_WORLD * sizeof(uint32_t), sizeof(w), cudaMemcpyDeviceToHost));
    *timed_out = (int)w;
  });
}

void hppb200_exchange_destroy(hppb200_exchange* ex) {
  if (!ex) return;
  cudaSetDevice(ex->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < ex->world; ++r)
    if (r != ex->rank && ex->opened[r] && ex->peer[r]) cudaIpcCloseMemHandle(ex->peer[r]);
  if (ex->d_peer) cudaFree(ex->d_peer);
  if (ex->d_ticket) cudaFree(ex->d_ticket);
  if (ex->local) cudaFree(ex->local);
  delete ex;
}

}  // extern "C"
