// Peer-memory plumbing for the fused Ulysses exchange (one process per GPU, NVLink / NVSwitch).
//
// The head<->sequence all-to-all around attention is not a separate collective here: K2 stores the
// normalised/rotated q, k (and v) head vectors straight into the destination rank's HBM, and the
// attention epilogue stores O straight into the token owner's HBM (qi_qknorm_rope_scatter,
// qi_attention_fwd_scatter).  What remains of the collective is a flag barrier between the
// producer kernel and the consumer kernel on the other GPUs, which is this file:
//
//   qi_peer_alloc / qi_peer_export / qi_peer_import   cudaMalloc + CUDA IPC handles (host layer
//                                                     exchanges the 64-byte handles over
//                                                     torch.distributed)
//   qi_peer_barrier                                   one 32-thread kernel: release-increment my slot
//                                                     in every peer's signal array, acquire-spin on
//                                                     my own array until every peer has arrived.
//
// The barrier is epoch based (monotone counters, no reset), lives entirely in device memory and is
// therefore CUDA-graph replayable.  A spin that exceeds the timeout sets an error flag instead of
// hanging the GPU; once the flag is set every later barrier returns immediately.
#include "qi_common.cuh"
#include "../../include/qi_b200.h"
#include <cstring>
#include <mutex>
#include <string>

namespace qi {

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// sig[p] = device pointer (peer mapped) to rank p's signal array [n] of u32.  state[0] = epoch of the
// last completed barrier, state[1] = sticky error flag.
__global__ void __launch_bounds__(32) peer_barrier_kernel(uint32_t* const* __restrict__ sig, int n,
                                                          int me, uint32_t* __restrict__ state,
                                                          unsigned long long timeout_ns) {
  const int p = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();   // the flag must not overtake the stores of the kernels before it
  const uint32_t epoch = state[0] + 1;
  const uint32_t failed = state[1];
  __syncwarp();
  if (p < n) {
    // everything this GPU wrote to peer memory in earlier kernels of the stream is ordered before
    // the flag by the system-scope fence + release
    __threadfence_system();
    uint32_t* remote = sig[p] + me;
    asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(remote) : "memory");
    if (!failed) {
      const uint32_t* mine = sig[me] + p;
      const unsigned long long t0 = global_ns();
      for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
        if (static_cast<int32_t>(v - epoch) >= 0) break;
        if (global_ns() - t0 > timeout_ns) {
          atomicExch(&state[1], 1u);
          break;
        }
        __nanosleep(32);
      }
    }
  }
  __syncwarp();
  if (p == 0) state[0] = epoch;
}

}  // namespace qi

using namespace qi;

extern "C" int qi_peer_alloc(void** ptr, unsigned long long bytes) {
  if (!ptr || bytes == 0) return QI_ERR_ARG;
  if (qi_num_sms() <= 0) return QI_ERR_CUDA;
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) return QI_ERR_CUDA;
  if (cudaMemset(p, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
    cudaFree(p);
    return QI_ERR_CUDA;
  }
  *ptr = p;
  return QI_OK;
}

extern "C" int qi_peer_free(void* ptr) {
  if (!ptr) return QI_OK;
  return cudaFree(ptr) == cudaSuccess ? QI_OK : QI_ERR_CUDA;
}

extern "C" int qi_peer_export(void* ptr, void* handle64 /* host, 64 bytes */) {
  if (!ptr || !handle64) return QI_ERR_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, ptr) != cudaSuccess) return QI_ERR_CUDA;
  memcpy(handle64, &h, sizeof(h));
  return QI_OK;
}

extern "C" int qi_peer_import(const void* handle64 /* host, 64 bytes */, void** ptr) {
  if (!handle64 || !ptr) return QI_ERR_ARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  void* p = nullptr;
  if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) return QI_ERR_CUDA;
  *ptr = p;
  return QI_OK;
}

extern "C" int qi_peer_unmap(void* ptr) {
  if (!ptr) return QI_OK;
  return cudaIpcCloseMemHandle(ptr) == cudaSuccess ? QI_OK : QI_ERR_CUDA;
}

extern "C" int qi_peer_barrier(void* const* signal_ptrs_dev, int n_peers, int my_index,
                               uint32_t* state_dev, unsigned int timeout_ms, cudaStream_t stream) {
  if (!signal_ptrs_dev || !state_dev || n_peers < 1 || n_peers > 8 || my_index < 0 ||
      my_index >= n_peers)
    return QI_ERR_ARG;
  if (qi_num_sms() <= 0) return QI_ERR_CUDA;
  const unsigned long long ns = (unsigned long long)(timeout_ms ? timeout_ms : 2000) * 1000000ull;
  if (qi_launch(peer_barrier_kernel, dim3(1), dim3(32), 0, stream,
                reinterpret_cast<uint32_t* const*>(signal_ptrs_dev), n_peers, my_index, state_dev,
                ns) != cudaSuccess)
    return QI_ERR_CUDA;
  return QI_OK;
}

// ------------------------------------------------------------------ qi_peer_barrier_static
// The same barrier kernel for callers that cannot obtain device memory for its bookkeeping.
//
// qi_peer_barrier needs two small pieces of device memory besides the peer-mapped signal arrays it signals
// through: the table of signal-array pointers the kernel indexes, and the {epoch, timeout flag} words.  Those are
// scratch, and the caller normally allocates them.  A serving worker whose HBM has been handed wholesale to the
// engine's block pool after memory profiling, a step loop that must not take cudaMalloc's device-wide
// synchronisation, or a C / Go host that binds this library and owns no device allocator has no way to do that.
// For them the scratch lives here, in a fixed array inside the library image: 256 bytes on every GPU that loads
// the library, held until the process exits.
//
//   bytes [0, 64)   pointer table, QI_PEER_BARRIER_STATIC_MAX_PEERS (8) entries
//   bytes [64, 72)  state: epoch of the last barrier, sticky timeout flag
//
// `g_barrier_scratch` below is only a handle in host code: the kernel must be given the address that
// cudaGetSymbolAddress reports for the device that is current, which is looked up once per device.
static constexpr int kStaticMaxPeers = QI_PEER_BARRIER_STATIC_MAX_PEERS;
static constexpr size_t kStaticTableBytes = kStaticMaxPeers * sizeof(void*);
static constexpr size_t kStaticBytes = 256;
static_assert(kStaticTableBytes + 2 * sizeof(uint32_t) <= kStaticBytes, "scratch layout");
static __device__ __align__(256) unsigned char g_barrier_scratch[kStaticBytes];

void qi_set_last_error(const char* msg);   // qi_api.cu

namespace qi {
struct StaticPeerTable {
  uint32_t* p[kStaticMaxPeers];
};
// Writes the caller's signal-array pointers into the table part of the scratch (device code may name the
// array directly).  The barrier kernel that follows in the stream reads them after griddepcontrol.wait.
__global__ void __launch_bounds__(32) peer_static_table_kernel(StaticPeerTable t, int n) {
  pdl_launch_dependents();
  pdl_wait();   // an earlier barrier of the stream may still be reading the previous table
  if (threadIdx.x < n) reinterpret_cast<uint32_t**>(g_barrier_scratch)[threadIdx.x] = t.p[threadIdx.x];
}
}  // namespace qi

static std::mutex g_static_mutex;
static unsigned char* g_static_addr[64];            // device address of g_barrier_scratch, per device ordinal
static void* g_static_table[64][kStaticMaxPeers];   // what the table on that device holds (host copy)

// Device address of the scratch on the current device (nullptr + error string on failure).  The first call on a
// device does the cudaGetSymbolAddress look-up, with that device current.
static unsigned char* static_scratch_for_current_device(int* dev_out) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    qi_set_last_error("qi_peer_barrier_static: no current CUDA device");
    return nullptr;
  }
  *dev_out = dev;
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (!g_static_addr[dev]) {
    unsigned char* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_barrier_scratch);
    if (e != cudaSuccess || !p) {
      qi_set_last_error((std::string("qi_peer_barrier_static: cudaGetSymbolAddress: ") +
                         cudaGetErrorString(e)).c_str());
      return nullptr;
    }
    g_static_addr[dev] = p;
  }
  return g_static_addr[dev];
}

extern "C" int qi_peer_barrier_static(void* const* signal_ptrs /* host */, int n_peers, int my_index,
                                      int restart, uint32_t* state_out_dev, unsigned int timeout_ms,
                                      cudaStream_t stream) {
  if (n_peers > kStaticMaxPeers) {
    qi_set_last_error(("qi_peer_barrier_static: n_peers = " + std::to_string(n_peers) +
                       " does not fit the library's fixed pointer table (" +
                       std::to_string(kStaticMaxPeers) + " entries)").c_str());
    return QI_ERR_ARG;
  }
  if (!signal_ptrs || n_peers < 1 || my_index < 0 || my_index >= n_peers) return QI_ERR_ARG;
  for (int p = 0; p < n_peers; ++p)
    if (!signal_ptrs[p]) return QI_ERR_ARG;
  if (qi_num_sms() <= 0) return QI_ERR_CUDA;
  int dev = 0;
  unsigned char* scratch = static_scratch_for_current_device(&dev);
  if (!scratch) return QI_ERR_CUDA;
  uint32_t* const* table = reinterpret_cast<uint32_t* const*>(scratch);
  uint32_t* state = reinterpret_cast<uint32_t*>(scratch + kStaticTableBytes);
  // The array still holds whatever the previous call left.  The epoch and the timeout flag carry over on
  // purpose, except when the caller starts a new count on freshly zeroed signal arrays; the table is rewritten
  // whenever it differs from what the previous call on this device stored.
  cudaError_t e = cudaSuccess;
  bool same;
  {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    same = memcmp(g_static_table[dev], signal_ptrs, n_peers * sizeof(void*)) == 0;
    if (!same) memcpy(g_static_table[dev], signal_ptrs, n_peers * sizeof(void*));
  }
  if (!same || restart) {
    StaticPeerTable t = {};
    for (int p = 0; p < n_peers; ++p) t.p[p] = static_cast<uint32_t*>(signal_ptrs[p]);
    e = qi_launch(peer_static_table_kernel, dim3(1), dim3(32), 0, stream, t, n_peers);
  }
  if (e == cudaSuccess && restart) e = cudaMemsetAsync(state, 0, 2 * sizeof(uint32_t), stream);
  const unsigned long long ns = (unsigned long long)(timeout_ms ? timeout_ms : 2000) * 1000000ull;
  if (e == cudaSuccess)
    e = qi_launch(peer_barrier_kernel, dim3(1), dim3(32), 0, stream, table, n_peers, my_index, state, ns);
  if (e == cudaSuccess && state_out_dev)
    e = cudaMemcpyAsync(state_out_dev, state, 2 * sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream);
  if (e != cudaSuccess) {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    memset(g_static_table[dev], 0, sizeof(g_static_table[dev]));   // unknown device contents: rewrite next time
    qi_set_last_error((std::string("qi_peer_barrier_static: ") + cudaGetErrorString(e)).c_str());
    return QI_ERR_CUDA;
  }
  return QI_OK;
}
