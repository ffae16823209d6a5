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
