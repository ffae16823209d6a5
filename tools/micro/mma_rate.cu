// Microbenchmark: cycles per tcgen05.mma (kind::f16, M = 128, K = 16, cta_group::1) by operand source and N.
//   mode 0/1/2: A and B from shared memory (SS), N = 128 / 64 / 256
//   mode 3/4/5: A from TMEM (TS),                N = 128 / 64 / 256
// One thread per CTA issues `groups` x 8 MMAs into one accumulator, commits, waits; clock64 around it.  Run on every
// SM at once (grid = SM count).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../vllm_omni_draft_b200/csrc
#include "qi_common.cuh"
#include <cstdio>
using namespace qi;

template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) rate_kernel(int groups, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (32768 + 65536) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 1 && elect_one()) {
    constexpr uint32_t idesc = make_idesc_bf16(128, N, 0, 0);
    const uint32_t a_addr = smem_u32(smem);            // A: 128 rows x 64 bf16 (16 KB), two of them
    const uint32_t b_addr = smem_u32(smem + 32768);    // B: up to 256 rows x 64 bf16, two sub-tiles
    const long long t0 = clock64();
    for (int g = 0; g < groups; ++g) {
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const uint64_t da = make_smem_desc(a_addr + s * 16384, 0, 1024);
        const uint64_t db = make_smem_desc(b_addr + s * 32768, 0, 1024);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (TS) mma_ts(tm + 256, tm + (s * 4 + k) * 8, db + 2 * k, idesc, 1u);
          else mma_ss(tm + 256, da + 2 * k, db + 2 * k, idesc, 1u);
        }
      }
    }
    tc_commit(&bar);
    mbar_wait(&bar, 0);
    const long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}

// The attention pipeline's step: 4 TS MMAs N=128 (P*V) + 8 TS MMAs N=64 (Q*K^T) = 512 clk of tensor work, with
// `commits` tcgen05.commit (to distinct mbarriers nobody waits on) spread behind the two groups, and optionally one
// wait per step on a barrier that the previous step's commit completes (a dependent round trip through the pipe).
template <int COMMITS, int WAITDIST>
__global__ void __launch_bounds__(128, 1) step_kernel(int steps, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __shared__ uint64_t bars[8];
  __shared__ uint64_t done;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (32768 + 65536) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1);
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 1 && elect_one()) {
    constexpr uint32_t idesc_qk = make_idesc_bf16(128, 64, 0, 0);
    constexpr uint32_t idesc_pv = make_idesc_bf16(128, 128, 0, 1);
    const uint32_t b_addr = smem_u32(smem + 32768);
    const long long t0 = clock64();
    for (int h = 0; h < steps; ++h) {
      if (WAITDIST > 0 && h >= WAITDIST) mbar_wait(&bars[0], (h - WAITDIST) & 1);   // completed by the commit of step h - WAITDIST
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const uint64_t db = make_smem_desc(b_addr + kk * 2048, 8192, 1024);
        mma_ts(tm + 256 + (h & 1) * 128, tm + 64 + (h % 3) * 64 + kk * 8, db, idesc_pv, 1u);
      }
      if (COMMITS >= 1) tc_commit(&bars[0]);
      if (COMMITS >= 3) tc_commit(&bars[1]);
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const uint64_t db = make_smem_desc(b_addr + 16384 + s * 8192, 0, 1024);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_ts(tm + 64 + (h % 3) * 64, tm + (s * 4 + k) * 8, db + 2 * k, idesc_qk, (s | k) ? 1u : 0u);
      }
      if (COMMITS >= 2) tc_commit(&bars[2]);
      if (COMMITS >= 4) tc_commit(&bars[3]);
    }
    tc_commit(&done);
    mbar_wait(&done, 0);
    out[blockIdx.x] = clock64() - t0;
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}

template <int COMMITS, int WAITDIST>
static void run_step(int steps, int sms) {
  long long* d;
  cudaMalloc(&d, sms * sizeof(long long));
  const int smem = 32768 + 65536 + 1024;
  cudaFuncSetAttribute(step_kernel<COMMITS, WAITDIST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  step_kernel<COMMITS, WAITDIST><<<sms, 128, smem>>>(steps, d);
  step_kernel<COMMITS, WAITDIST><<<sms, 128, smem>>>(steps, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[256];
  cudaMemcpy(h, d, sms * sizeof(long long), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("step (4 TS N=128 + 8 TS N=64 = 512 clk), %d commits/step, wait on the commit of %d steps back: %7.1f clk per step [%s]\n",
         COMMITS, WAITDIST, (double)mx / steps, cudaGetErrorString(e));
  cudaFree(d);
}

// The same step issued by the leader of a CTA pair: cta_group::2, M = 256, A from each CTA's TMEM, each CTA's shared
// memory supplies half of B (32 of the 64 keys for Q*K^T, 64 of the 128 head-dim columns for P*V).
__device__ __forceinline__ void mma_ts_pair_(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d_tmem),
               "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}
template <int MODE>   // 0: the attention step, 1: Q*K^T only (8 x N=64), 2: P*V only (4 x N=128)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) pair_step_kernel(int steps, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __shared__ uint64_t done;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (32768 + 65536) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc_pair<512>(&slot);
  fence_proxy_async();
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tm = slot;
  if (cluster_ctarank() == 0 && warp == 1 && elect_one()) {
    constexpr uint32_t idesc_qk = make_idesc_bf16(256, 64, 0, 0);
    constexpr uint32_t idesc_pv = make_idesc_bf16(256, 128, 0, 1);
    const uint32_t b_addr = smem_u32(smem + 32768);
    const long long t0 = clock64();
    for (int h = 0; h < steps; ++h) {
      if (MODE != 1) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t db = make_smem_desc(b_addr + kk * 2048, 8192, 1024);
          mma_ts_pair_(tm + 256 + (h & 1) * 128, tm + 64 + (h % 3) * 64 + kk * 8, db, idesc_pv, 1u);
        }
      }
      if (MODE != 2) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const uint64_t db = make_smem_desc(b_addr + 16384 + s * 4096, 0, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_ts_pair_(tm + 64 + (h % 3) * 64, tm + (s * 4 + k) * 8, db + 2 * k, idesc_qk, (s | k) ? 1u : 0u);
        }
      }
    }
    tc_commit_pair(&done, 1);
    mbar_wait(&done, 0);
    out[blockIdx.x >> 1] = clock64() - t0;
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc_pair<512>(tm);
}
template <int MODE>
static void run_pair(const char* what, int steps, int sms) {
  long long* d;
  cudaMalloc(&d, sms * sizeof(long long));
  const int smem = 32768 + 65536 + 1024;
  cudaFuncSetAttribute(pair_step_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  pair_step_kernel<MODE><<<sms / 2 * 2, 128, smem>>>(steps, d);
  pair_step_kernel<MODE><<<sms / 2 * 2, 128, smem>>>(steps, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[256];
  cudaMemcpy(h, d, (sms / 2) * sizeof(long long), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < sms / 2; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("CTA pair (cta_group::2, M=256, TS), %s: %7.1f clk per step [%s]\n", what, (double)mx / steps, cudaGetErrorString(e));
  cudaFree(d);
}

// MUFU.EX2 throughput per SM: f32 (one value per lane and instruction) against packed f16x2 (two).
template <int MODE>   // 0: ex2.approx.ftz.f32, 1: ex2.approx.f16x2
__global__ void __launch_bounds__(512, 1) ex2_kernel(int iters, long long* out, uint32_t* sink) {
  uint32_t v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = MODE ? 0xb800b400u + threadIdx.x : __float_as_uint(-0.5f - 0.001f * i);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(v[i]));
      else asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(v[i]));
    }
  }
  const long long t1 = clock64();
  uint32_t x = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) x ^= v[i];
  if (x == 0x12345678u) sink[0] = x;
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}
template <int MODE>
static void run_ex2(const char* name, int sms) {
  long long* d;
  uint32_t* sink;
  cudaMalloc(&d, sms * sizeof(long long));
  cudaMalloc(&sink, 4);
  const int iters = 4000;
  ex2_kernel<MODE><<<sms, 512>>>(iters, d, sink);
  ex2_kernel<MODE><<<sms, 512>>>(iters, d, sink);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[256];
  cudaMemcpy(h, d, sms * sizeof(long long), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
  const double instr = (double)iters * 8 * 16;   // warp instructions per SM (16 warps)
  printf("%-22s: %6.2f clk per warp instruction per SM sub-partition, %5.1f values per clk per SM [%s]\n", name,
         (double)mx / (instr / 4), instr * 32 * (MODE ? 2 : 1) / (double)mx, cudaGetErrorString(e));
  cudaFree(d);
  cudaFree(sink);
}

// Does kind::f16 accept A = f16 (in TMEM) with B = bf16 (in shared memory)?  A = 0.5 everywhere, B = 2.0 everywhere,
// one M = 128, N = 64, K = 16 MMA: every element of D must be 16.
__global__ void __launch_bounds__(128, 1) mixed_kernel(float* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 16384 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x40004000u;   // bf16 2.0
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
  uint32_t a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = 0x38003800u;   // f16 0.5 | 0.5
  tmem_st16(tm + lane_off, a);
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1 && elect_one()) {
    // c_format F32 (bit 4), a_format F16 (0 at bit 7), b_format BF16 (1 at bit 10), N = 64, M = 128
    const uint32_t idesc = (1u << 4) | (0u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t db = make_smem_desc(smem_u32(smem), 0, 1024);
    mma_ts(tm + 64, tm, db, idesc, 0u);
    tc_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  uint32_t r[16];
  tmem_ld16(tm + lane_off + 64, r);
  tmem_ld_wait();
  out[threadIdx.x] = __uint_as_float(r[0]) + __uint_as_float(r[15]);
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}
static void run_mixed() {
  float* d;
  cudaMalloc(&d, 128 * sizeof(float));
  cudaFuncSetAttribute(mixed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  mixed_kernel<<<1, 128, 32768>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  float h[128];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("kind::f16 with A = f16 (TMEM) x B = bf16 (smem): D[0][0] + D[0][15] = %g, row 127: %g (32 expected if the formats are honoured) [%s]\n",
         h[0], h[127], cudaGetErrorString(e));
  cudaFree(d);
}

template <int N, bool TS>
static void run(const char* name, int groups, int sms) {
  long long* d;
  cudaMalloc(&d, sms * sizeof(long long));
  const int smem = 32768 + 65536 + 1024;
  cudaFuncSetAttribute(rate_kernel<N, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int grid : {1, sms}) {
    rate_kernel<N, TS><<<grid, 128, smem>>>(groups, d);
    rate_kernel<N, TS><<<grid, 128, smem>>>(groups, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[256];
    cudaMemcpy(h, d, grid * sizeof(long long), cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("%-10s grid %3d: %7.1f clk per MMA (M=128 N=%d K=16)  -> %.0f%% of the 128*N/256 floor  [%s]\n", name, grid,
           (double)mx / (groups * 8), N, 100.0 * (128.0 * N / 256.0) / ((double)mx / (groups * 8)), cudaGetErrorString(e));
  }
  cudaFree(d);
}

int main() {
  setvbuf(stdout, nullptr, _IONBF, 0);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int groups = 2000;
  run_ex2<0>("ex2.approx.ftz.f32", sms);
  run_ex2<1>("ex2.approx.f16x2", sms);
  run<128, false>("SS N=128", groups, sms);
  run<64, false>("SS N=64", groups, sms);
  run<256, false>("SS N=256", groups, sms);
  run<128, true>("TS N=128", groups, sms);
  run<64, true>("TS N=64", groups, sms);
  run<256, true>("TS N=256", groups, sms);
  run_step<0, 0>(2000, sms);
  run_step<1, 0>(2000, sms);
  run_step<2, 0>(2000, sms);
  run_step<4, 0>(2000, sms);
  run_step<4, 1>(2000, sms);
  run_pair<0>("4 x N=128 + 8 x N=64 (512 clk at the single-CTA rate)", 2000, sms);
  run_pair<1>("8 x N=64 (256 clk)", 2000, sms);
  run_pair<2>("4 x N=128 (256 clk)", 2000, sms);
  run_mixed();   // last: an illegal instruction (the answer on sm_100a) poisons the context
  return 0;
}
