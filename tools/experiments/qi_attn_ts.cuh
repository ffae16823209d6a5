// attention_ts_kernel: the second attention kernel (variants 40+ / 56+ of qi_attention_set_variant) — one 128-query
// tile per CTA with Q resident in TMEM, K/V pipelined in 64-key halves.  Included by qi_attn.cu after the shared
// definitions (AttParams, exp2 helpers, pair_sync); measured slower than the default two-tile kernel on the
// benchmarked shapes and kept as a tested alternative.
#pragma once

namespace qi {

// ---------------------------------------------------------------------------------------------------------------
// TS: one 128-query tile per CTA, Q resident in TMEM, software-pipelined over 64-key HALVES of K/V.
//
// In the kernel above one Q tile's work is a dependent chain per K/V tile — S ready -> softmax (1 430 clk) -> second
// half of P*V -> next Q*K^T -> S ready — because S_{j+1} can only be written once P_j (which lives in S_j's columns)
// has been consumed: TMEM holds exactly two S tiles and two O tiles.  The two Q tiles hide part of each other's
// chain, not all of it (measured: MMA warp waits 30 % of its time for P, softmax warps 40 % for S).  Splitting S into
// 64-column halves breaks the chain, but an SS-mode Q*K^T with N = 64 takes 48 clk against 32 for the same MMA with A
// in TMEM (tools/micro/mma_rate.cu: A is re-read from shared memory for every MMA).  So here:
//   * Q (128 x DH bf16 = DH/2 TMEM columns) is the TMEM-resident A operand of Q*K^T: N = 64 MMAs run at the full
//     rate and shared memory only feeds B (K or V).
//   * S lives in three 64-column regions used round robin by successive halves h; P_h is written over the first 32
//     columns of its region.  The MMA thread issues P_h*V_h and then Q*K_{h+3}^T into the region P_h just left.
//   * Two softmax warpgroups take alternate halves of the SAME 128 rows (even halves / odd halves), each with its
//     own running reference, row sum and O accumulator (O_a, O_b); the two partial results are merged once, in
//     the epilogue.  A warpgroup's next S is produced while it still works on the current one, so it never waits
//     for more than the tensor pipe's own work (512 clk per half against ~720 clk of softmax on two warpgroups).
//   * K/V halves stream through deep TMA rings.  One Q tile per CTA doubles the K/V bytes per FLOP that an SM pulls
//     from L2, and that — not the tensor pipe — then bounds the kernel (measured: 1 180 clk per half with the K/V
//     waits, 600 without; TMA multicast in clusters of 2 does not reduce it).  PAIR: the two CTAs of a cluster
//     (adjacent Q tiles of one head) run cta_group::2 MMAs, M = 256: each CTA stages only HALF of every K/V half
//     (32 keys of K, 64 of the 128 head-dim columns of V) and the tensor cores read both halves, so an SM ingests the
//     same bytes per FLOP as in the two-tile kernel.  The leader CTA's MMA thread issues for both; both CTAs'
//     softmax warps arrive on the leader's P barriers; commits are multicast to both CTAs.
//   TMEM columns (per CTA): Q [0,64)  S regions [64,256)  O_a [256,384)  O_b [384,512).
template <int DH, bool PAIR>
struct TsCfg {
  static constexpr int HALF = 64;                              // keys per pipeline step
  static constexpr int NCTA = PAIR ? 2 : 1;
  static constexpr int NSUB = DH / 64;
  static constexpr int K_ROWS = HALF / NCTA;                   // key rows of a K half staged by one CTA
  static constexpr int K_SUB = K_ROWS * 128;                   // [K_ROWS keys x 64 bf16] swizzled sub-tile
  static constexpr int K_STAGE = K_SUB * NSUB;
  static constexpr int V_SUB = HALF * 128;                     // [64 keys x 64 head-dim columns]
  static constexpr int V_STAGE = V_SUB * (PAIR ? 1 : NSUB);    // PAIR: one CTA holds DH/2 = 64 columns of V
  static constexpr int NST = 6;                                // ring depth in halves, K and V each (the MMA thread's
                                                               // loop is unrolled over lcm(3 regions, 2 warpgroups, NST) = 6 steps)
  static constexpr int K_OFF = 0;
  static constexpr int V_OFF = NST * K_STAGE;
  static constexpr int BAR_OFF = V_OFF + NST * V_STAGE;
  static constexpr int XCH_OFF = BAR_OFF + 512;                // float [2 warpgroups][128 rows][2] = (m, l)
  static constexpr int SMEM_BYTES = XCH_OFF + 2048 + 1024;
  static constexpr int Q_COL = 0, S_COL = 64, O_COL = 256;
  static_assert(!PAIR || DH == 128, "the pair kernel splits V by 64-column sub-tiles");
  static_assert(NST * K_STAGE >= ATT_BM * DH * 2, "the epilogue stages O through the K ring");
  static_assert((4 * NST + 14) * 8 + 8 <= 512, "barrier block");
};

// D[tmem of both CTAs] (+)= A[tmem, 128 rows per CTA] * B[smem, N/2 rows per CTA]; leader CTA only
__device__ __forceinline__ void mma_ts_pair(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// mbarrier wait of the pipelined kernel: a protocol error (or a lost TMA) traps after ~2 s instead of hanging the GPU.
__device__ __forceinline__ void pipe_wait(uint64_t* bar, uint32_t parity) {
  uint32_t n = 0;
  unsigned long long t0 = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++n & 0xffffu) == 0) {
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 2000000000ull) __trap();
    }
  }
}

// Arrival on a barrier of this CTA, or (PAIR) on the same barrier of the pair's leader CTA.
template <bool PAIR>
__device__ __forceinline__ void ts_arrive(uint64_t* bar) {
  if (PAIR) mbar_arrive_cluster(mapa_shared(smem_u32(bar), 0));
  else mbar_arrive(bar);
}

template <int DH, bool MASK, int EMU, bool PAIR>
__device__ __forceinline__ void softmax_pipe_half(uint32_t s_addr, uint32_t o_addr, int kv_valid, int h,
                                                  float scale_log2, float& m_ref, float& l_sum, uint64_t* p_full,
                                                  uint64_t* pv_done, int lane, unsigned long long* pc = nullptr) {
  uint32_t r[2][32];
  long long c_prev = pc ? clock64() : 0;
  auto phase = [&](int i) {   // tuning aid: cycles per phase (one thread of the CTA)
    if (pc) {
      const long long now = clock64();
      pc[i] += now - c_prev;
      c_prev = now;
    }
  };
  tmem_ld32(s_addr, r[0]);
  tmem_ld32(s_addr + 32, r[1]);
  tmem_ld_wait();
  phase(0);
  if (MASK) {
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (c * 32 + i >= kv_valid) r[c][i] = 0xff800000u;   // -inf: contributes 0
  }
  float mxs[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) mxs[q] = -INFINITY;
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int i = 0; i < 16; ++i)
      mxs[i & 7] = fmaxf(fmaxf(mxs[i & 7], __uint_as_float(r[c][2 * i])), __uint_as_float(r[c][2 * i + 1]));
  const float mx = fmaxf(fmaxf(fmaxf(mxs[0], mxs[1]), fmaxf(mxs[2], mxs[3])),
                         fmaxf(fmaxf(mxs[4], mxs[5]), fmaxf(mxs[6], mxs[7])));
  const float m_new = fmaxf(m_ref, mx * scale_log2);
  const bool need = (m_new - m_ref) > 8.0f;   // also true on the first half (m_ref = -inf)
  phase(1);
  if (__any_sync(0xffffffffu, need)) {
    const float f = need ? fast_exp2(m_ref - m_new) : 1.0f;   // exp2(-inf) = 0 on half 0
    if (need) {
      l_sum *= f;
      m_ref = m_new;
    }
    if (h > 0) {
      pipe_wait(pv_done, (h - 1) & 1);   // O must hold every P*V issued so far before it is scaled
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < DH / 32; ++c) {
        uint32_t o[32];
        tmem_ld32(o_addr + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * f);
        tmem_st32(o_addr + c * 32, o);
      }
    }
  }
  phase(2);
  constexpr int NE = EMU & 15;
  constexpr int DEG = (EMU & 16) ? 2 : 3;
  const float2 sc2 = make_float2(scale_log2, scale_log2);
  const float2 nm2 = make_float2(-m_ref, -m_ref);
  float2 acc[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) acc[q] = make_float2(0.f, 0.f);
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    uint32_t pk[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float2 x = __ffma2_rn(make_float2(__uint_as_float(r[c][2 * i]), __uint_as_float(r[c][2 * i + 1])),
                                  sc2, nm2);
      float2 p;
      if (((i % 8) * NE) % 8 < NE) {
        p = exp2_poly2<DEG>(x);
      } else {
        p.x = fast_exp2(x.x);
        p.y = fast_exp2(x.y);
      }
      acc[i & 3] = __fadd2_rn(acc[i & 3], p);
      pk[i] = pack_bf16(p.x, p.y);
    }
    tmem_st16(s_addr + c * 16, pk);   // P_h: packed bf16 over the first 32 columns of this region
  }
  const float2 a01 = __fadd2_rn(acc[0], acc[1]), a23 = __fadd2_rn(acc[2], acc[3]);
  l_sum += (a01.x + a01.y) + (a23.x + a23.y);
  phase(3);
  tmem_st_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) ts_arrive<PAIR>(p_full);
  phase(4);
}

// tuning aid: timeline of CTA (0,0,0) — dbg[16 + 8*h + i] = clock64 at event i of half h (see tools/attn_ts_diag.py)
__device__ __forceinline__ void ts_trace(unsigned long long* dbg, bool on, int h, int i) {
  if (on && h < 96) dbg[16 + 8 * h + i] = clock64();
}

template <int DH, int EMU, bool PAIR>
__global__ void __cluster_dims__(PAIR ? 2 : 1, 1, 1) __launch_bounds__(ATT_THREADS, 1)
attention_ts_kernel(const __grid_constant__ AttParams P) {
  using Cfg = TsCfg<DH, PAIR>;
  constexpr int W_K = 8, W_MMA = 9, W_V = 10, W_QK = 11;   // third warpgroup: K loader, P*V issuer (TMEM owner), V loader, Q*K^T issuer
  constexpr int HALF = Cfg::HALF, NST = Cfg::NST;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::BAR_OFF);
  uint64_t* k_full = bars;                  // [NST]  (PAIR: the leader's collect both CTAs' TMA bytes)
  uint64_t* k_empty = bars + NST;           // [NST]
  uint64_t* v_full = bars + 2 * NST;        // [NST]
  uint64_t* v_empty = bars + 3 * NST;       // [NST]
  uint64_t* s_full = bars + 4 * NST;        // [3 regions]
  uint64_t* p_full = s_full + 3;            // [3 regions]  (PAIR: the leader's collect both CTAs' softmax warps)
  uint64_t* pv_done = p_full + 3;           // [2 warpgroups], one phase per P*V into that warpgroup's O
  uint64_t* o_done = pv_done + 2;
  uint64_t* q_ready = o_done + 1;
  uint64_t* r_free = q_ready + 1;           // [3 regions]  P*V of the half that occupied the region has completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(r_free + 3);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * ATT_BM;
  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int n_half = (P.seq_kv + HALF - 1) / HALF;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0;
  const bool trace = P.dbg != nullptr && (blockIdx.x | blockIdx.y | blockIdx.z) == 0;
  pdl_launch_dependents();

  if (warp == W_K && lane == 0) {
    tma_prefetch_desc(&P.tmap_k);
    tma_prefetch_desc(&P.tmap_v);
    for (int i = 0; i < NST; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 3; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4 * Cfg::NCTA);
      mbar_init(&r_free[i], 1);
    }
    mbar_init(&pv_done[0], 1);
    mbar_init(&pv_done[1], 1);
    mbar_init(o_done, 1);
    mbar_init(q_ready, 8 * Cfg::NCTA);
    fence_barrier_init();
  }
  if (warp == W_MMA) {
    if (PAIR) tmem_alloc_pair<512>(tmem_slot);
    else tmem_alloc<512>(tmem_slot);
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();   // both CTAs' barriers and TMEM exist before anything is signalled across the pair
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == W_K || warp == W_V) {
    // ------------------------------------------------------------------ TMA producers: K ring, V ring
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    const bool is_k = warp == W_K;
    if (is_k) peer_wait_all_warp(P.wait, lane);
    if (elect_one()) {
      uint64_t* full = is_k ? k_full : v_full;
      uint64_t* empty = is_k ? k_empty : v_empty;
      const CUtensorMap* map = is_k ? &P.tmap_k : &P.tmap_v;
      uint8_t* ring = smem + (is_k ? Cfg::K_OFF : Cfg::V_OFF);
      const int stage_bytes = is_k ? Cfg::K_STAGE : Cfg::V_STAGE;
      int st = 0;
      uint32_t ph = 0;
      for (int h = 0; h < n_half; ++h, ++st) {
        if (st == NST) st = 0, ph ^= 1;
        pipe_wait(&empty[st], ph ^ 1);
        ts_trace(P.dbg, trace, h, is_k ? 0 : 1);
        uint8_t* dst = ring + st * stage_bytes;
        if (PAIR) {
          // this CTA's half of the half: 32 of the 64 keys of K, 64 of the 128 head-dim columns of V; the bytes of
          // both CTAs complete on the leader's barrier
          if (rank == 0) mbar_expect_tx(&full[st], 2 * stage_bytes);
          const uint32_t bar = mapa_shared(smem_u32(&full[st]), 0);
          if (is_k) {
            for (int s = 0; s < Cfg::NSUB; ++s)
              tma_load_4d_pair(dst + s * Cfg::K_SUB, map, bar, s * 64, head, h * HALF + (int)rank * Cfg::K_ROWS, b);
          } else {
            tma_load_4d_pair(dst, map, bar, (int)rank * 64, head, h * HALF, b);
          }
        } else {
          mbar_expect_tx(&full[st], stage_bytes);
          for (int s = 0; s < Cfg::NSUB; ++s)
            tma_load_4d(dst + s * (is_k ? Cfg::K_SUB : Cfg::V_SUB), map, &full[st], s * 64, head, h * HALF, b);
        }
      }
    }
    __syncwarp();
  } else if (warp == W_MMA || warp == W_QK) {
    // ------------------------------------------------------------------ MMA issuers (PAIR: the leader CTA's only)
    // One thread could not issue a step's 12 MMAs, 4 commits and 3 barrier waits in the 512 clk the tensor pipe
    // needs for them: its scalar instruction stream ran at 10-15 clk per instruction (ncu source page, in-kernel
    // timeline: 1 200-1 450 clk per step with the tensor pipe 33 % active and every barrier already open).  So P*V
    // and Q*K^T are issued by two threads in two warps (two SM sub-partitions).  They are ordered through barriers
    // only: Q*K_{h+3}^T, which overwrites the region of P_h, waits for `r_free` (committed behind P_h*V_h).  Both
    // loops are unrolled over 6 steps = lcm(3 S regions, 2 O accumulators, 6 ring stages), so that every TMEM
    // column, descriptor and barrier is a base plus a compile-time constant.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    static_assert(NST == 6, "unrolled over 6 steps");
    if (rank == 0 && elect_one()) {
      const uint32_t s_col = tmem_base + Cfg::S_COL;
      auto mma = [&](uint32_t d, uint32_t a, uint64_t bd, uint32_t idesc, uint32_t acc) {
        if (PAIR) mma_ts_pair(d, a, bd, idesc, acc);
        else mma_ts(d, a, bd, idesc, acc);
      };
      auto commit = [&](uint64_t* bar) {
        if (PAIR) tc_commit_pair(bar, 3);
        else tc_commit(bar);
      };
      // Descriptors of ring stage st = those of stage 0 + st * (stage bytes >> 4) in the 14-bit address field
      // (shared-memory addresses are below 256 KB, so the field cannot carry).
      const long long t_begin = clock64();
      if (warp == W_QK) {
        // S[region] = Q * K_h^T : A = Q from TMEM (8 packed columns per 16 dims), B = the K half in a ring stage
        constexpr uint32_t idesc_qk = make_idesc_bf16(ATT_BM * Cfg::NCTA, HALF, 0, 0);
        constexpr uint32_t KSTEP = Cfg::K_STAGE >> 4;
        const uint32_t q_col = tmem_base + Cfg::Q_COL;
        const uint64_t kdesc0 = make_smem_desc(smem_u32(smem + Cfg::K_OFF), 0, 1024);
        mbar_wait(q_ready, 0);
        tc_fence_after();
        uint32_t bp = 0;                      // parity of the ring barriers during this pass over the 6 stages
        for (int h0 = 0; h0 < n_half; h0 += 6, bp ^= 1) {
#pragma unroll
          for (int u = 0; u < 6; ++u) {
            const int h = h0 + u;
            if (h >= n_half) break;
            const int r = u % 3;
            // the region's previous tenant is half h - 3: its P*V is use number (h - 3) / 3 of r_free[r]
            if (h >= 3) mbar_wait(&r_free[r], u >= 3 ? 0u : 1u);   // ((h - 3) / 3) & 1 with h0 / 3 even
            mbar_wait(&k_full[u], bp);
            tc_fence_after();
            ts_trace(P.dbg, trace, h, 2);
            const uint64_t kd = kdesc0 + u * KSTEP;
#pragma unroll
            for (int s = 0; s < Cfg::NSUB; ++s)
#pragma unroll
              for (int k = 0; k < 4; ++k)
                mma(s_col + r * HALF, q_col + (s * 4 + k) * 8, kd + s * (Cfg::K_SUB >> 4) + 2 * k, idesc_qk,
                    (s | k) != 0 ? 1u : 0u);
            commit(&s_full[r]);
            commit(&k_empty[u]);
          }
        }
      } else {
        // O[wg] += P[region] * V_h : 16 keys per MMA = 8 packed columns of P and 2048 B of V; one CTA: the second 64
        // columns of the head dim live one sub-tile (LBO) further; PAIR: they live in the other CTA
        constexpr uint32_t idesc_pv = make_idesc_bf16(ATT_BM * Cfg::NCTA, DH, 0, 1);   // B (=V) is MN-major
        constexpr uint32_t VSTEP = Cfg::V_STAGE >> 4;
        const uint32_t o_col = tmem_base + Cfg::O_COL;
        const uint64_t vdesc0 = make_smem_desc(smem_u32(smem + Cfg::V_OFF), Cfg::V_SUB, 1024);
        uint32_t bp = 0;
        for (int h0 = 0; h0 < n_half; h0 += 6, bp ^= 1) {
#pragma unroll
          for (int u = 0; u < 6; ++u) {
            const int h = h0 + u;
            if (h >= n_half) break;
            const int r = u % 3;
            mbar_wait(&v_full[u], bp);
            if (P.ts_diag < 3) mbar_wait(&p_full[r], (u / 3) & 1);   // (h / 3) & 1 with h0 / 3 even
            tc_fence_after();
            ts_trace(P.dbg, trace, h, 4);
            const uint64_t vd = vdesc0 + u * VSTEP;
#pragma unroll
            for (int kk = 0; kk < HALF / 16; ++kk)
              mma(o_col + (u & 1) * 128, s_col + r * HALF + kk * 8, vd + kk * (2048 >> 4), idesc_pv,
                  (h >= 2 || kk != 0) ? 1u : 0u);
            commit(&pv_done[u & 1]);
            commit(&v_empty[u]);
            commit(&r_free[r]);
            ts_trace(P.dbg, trace, h, 5);
          }
        }
        commit(o_done);
        if (P.dbg) atomicAdd(&P.dbg[0], (unsigned long long)(clock64() - t_begin));
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax warpgroups: wg 0 even halves, wg 1 odd
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int wg = warp >> 2;
    const int quad = warp & 3;               // TMEM lane quadrant
    const int row_l = quad * 32 + lane;      // this thread's query row inside the tile
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const uint32_t s_addr = tmem_base + lane_off + Cfg::S_COL;
    const uint32_t o_addr = tmem_base + lane_off + Cfg::O_COL + wg * 128;
    {   // Q row -> TMEM: this warpgroup's half of the head dim, straight from global memory (once per CTA)
      constexpr int NQ = DH / 4;             // packed 32-bit columns per thread
      uint32_t qv[NQ];
      const int row = q0 + row_l;
      if (row < P.seq_q) {
        const uint4* src = reinterpret_cast<const uint4*>(P.q + (size_t)b * P.q_bs + (size_t)row * P.ld_q +
                                                          (size_t)head * P.q_hs + wg * (DH / 2));
#pragma unroll
        for (int i = 0; i < NQ / 4; ++i) {
          const uint4 u = __ldg(src + i);
          qv[4 * i] = u.x, qv[4 * i + 1] = u.y, qv[4 * i + 2] = u.z, qv[4 * i + 3] = u.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < NQ; ++i) qv[i] = 0u;
      }
      if constexpr (NQ == 32) tmem_st32(tmem_base + lane_off + Cfg::Q_COL + wg * NQ, qv);
      else tmem_st16(tmem_base + lane_off + Cfg::Q_COL + wg * NQ, qv);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) ts_arrive<PAIR>(q_ready);
    }
    float m_ref = -INFINITY;   // reference max (scaled, log2 domain) of this warpgroup's halves
    float l_sum = 0.f;
    const int n_full = P.seq_kv / HALF;      // halves with all 64 keys valid
    const bool prof = P.dbg != nullptr && (threadIdx.x & 127) == 0 && wg == 0 && rank == 0;
    unsigned long long w_s = 0;
    unsigned long long pcs[5] = {0, 0, 0, 0, 0};   // S load, row max, O rescale, P (exp2 + pack + store), publish
    const long long t_begin = clock64();
    int region = wg;               // h % 3 and (h / 3) & 1 of this warpgroup's halves h = wg, wg + 2, ...
    uint32_t rpar = 0;
    for (int h = wg; h < n_half; h += 2, region += 2) {
      if (region >= 3) region -= 3, rpar ^= 1;
      const long long c0 = prof ? clock64() : 0;
      pipe_wait(&s_full[region], rpar);
      if (prof) w_s += clock64() - c0;
      ts_trace(P.dbg, trace && (threadIdx.x & 127) == 0, h, 6);
      tc_fence_after();
      if (P.ts_diag) {   // timing experiments only: the result is garbage
        if (P.ts_diag == 2) {
          uint32_t r[2][32];
          tmem_ld32(s_addr + region * HALF, r[0]);
          tmem_ld32(s_addr + region * HALF + 32, r[1]);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) pk[i] = r[0][i] ^ r[1][i];
          tmem_st16(s_addr + region * HALF, pk);
          tmem_st16(s_addr + region * HALF + 16, pk);
          tmem_st_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) ts_arrive<PAIR>(&p_full[region]);
        l_sum = 1.f;
        m_ref = 0.f;
        continue;
      }
      if (h < n_full) {
        softmax_pipe_half<DH, false, EMU, PAIR>(s_addr + region * HALF, o_addr, HALF, h >> 1, P.scale_log2, m_ref,
                                                l_sum, &p_full[region], &pv_done[wg], lane, prof ? pcs : nullptr);
      } else {
        softmax_pipe_half<DH, true, EMU, PAIR>(s_addr + region * HALF, o_addr, P.seq_kv - h * HALF, h >> 1,
                                               P.scale_log2, m_ref, l_sum, &p_full[region], &pv_done[wg], lane);
      }
      ts_trace(P.dbg, trace && (threadIdx.x & 127) == 0, h, 7);
    }
    if (prof) {
      atomicAdd(&P.dbg[8], (unsigned long long)(clock64() - t_begin));
      atomicAdd(&P.dbg[9], w_s);
      for (int i = 0; i < 5; ++i) atomicAdd(&P.dbg[11 + i], pcs[i]);
    }
    // merge the two warpgroups' partial softmaxes: O = (O_a*2^(m_a-m) + O_b*2^(m_b-m)) / (l_a*2^(m_a-m) + l_b*2^(m_b-m))
    float* xch = reinterpret_cast<float*>(smem + Cfg::XCH_OFF);
    xch[(wg * 128 + row_l) * 2] = m_ref;
    xch[(wg * 128 + row_l) * 2 + 1] = l_sum;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    const float m_o = xch[((wg ^ 1) * 128 + row_l) * 2];
    const float l_o = xch[((wg ^ 1) * 128 + row_l) * 2 + 1];
    const bool has_b = n_half > 1;           // with a single half warpgroup 1 never ran: O_b is uninitialised
    const float m_all = fmaxf(m_ref, m_o);
    const float a_me = fast_exp2(m_ref - m_all);      // exp2(-inf) = 0 for a warpgroup that saw no keys
    const float a_ot = fast_exp2(m_o - m_all);
    const float inv = 1.f / (l_sum * a_me + l_o * a_ot);
    const float al_a = (wg == 0 ? a_me : a_ot) * inv;
    const float al_b = (wg == 0 ? a_ot : a_me) * inv;
    pipe_wait(o_done, 0);
    tc_fence_after();
    // this warpgroup converts its half of the head-dim columns (of O_a and O_b); rows are staged through the K ring
    // (all K/V halves have been consumed) so that global stores are DH*2 contiguous bytes per DH/8 lanes
    constexpr int PPR = DH / 8;          // 16-byte pieces per row
    constexpr int RPI = 32 / PPR;        // rows per store instruction
    const uint32_t oa = tmem_base + lane_off + Cfg::O_COL + wg * (DH / 2);
    uint8_t* stage = smem + Cfg::K_OFF + quad * (32 * DH * 2);
#pragma unroll 1
    for (int c = 0; c < DH / 64; ++c) {
      uint32_t ra[32], rb[32];
      tmem_ld32(oa + c * 32, ra);
      if (has_b) tmem_ld32(oa + 128 + c * 32, rb);
      tmem_ld_wait();
      if (!has_b) {
#pragma unroll
        for (int i = 0; i < 32; ++i) rb[i] = 0u;
      }
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
          v[i] = __uint_as_float(ra[g * 8 + i]) * al_a + __uint_as_float(rb[g * 8 + i]) * al_b;
        uint4 o;
        o.x = pack_bf16(v[0], v[1]);
        o.y = pack_bf16(v[2], v[3]);
        o.z = pack_bf16(v[4], v[5]);
        o.w = pack_bf16(v[6], v[7]);
        const int piece = wg * (DH / 16) + c * 4 + g;
        *reinterpret_cast<uint4*>(stage + lane * (DH * 2) + ((piece ^ (lane & 7)) << 4)) = o;
      }
    }
    pair_sync(2 + quad);                     // the two warps of this lane quadrant have staged their columns
    const int row_base = q0 + quad * 32;
    constexpr int ITS = 32 / RPI / 2;        // each of the two warps stores half of the 32 rows
#pragma unroll 4
    for (int it = wg * ITS; it < (wg + 1) * ITS; ++it) {
      const int rr = it * RPI + lane / PPR;
      const int piece = lane % PPR;
      const int row = row_base + rr;
      const uint4 o = *reinterpret_cast<const uint4*>(stage + rr * (DH * 2) + ((piece ^ (rr & 7)) << 4));
      if (row < P.seq_q) {
        __nv_bfloat16* orow;
        if (P.n_peer > 0) {
          int p = 0;
          while (p + 1 < P.n_peer && row >= P.o_prefix[p + 1]) ++p;
          orow = P.o_peer[p] + P.o_chunk_off + (size_t)b * P.o_bs + (size_t)(row - P.o_prefix[p]) * P.ld_o +
                 (size_t)head * P.o_hs;
        } else {
          orow = P.o + (size_t)b * P.o_bs + (size_t)row * P.ld_o + (size_t)head * P.o_hs;
        }
        *reinterpret_cast<uint4*>(orow + piece * 8) = o;
      }
    }
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all();   // no CTA leaves while the pair's MMAs or arrivals may still touch it
  else __syncthreads();
  if (threadIdx.x == 0) peer_notify_last_cta(P.notify, gridDim.x * gridDim.y * gridDim.z);
  if (warp == W_MMA) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair<512>(tmem_base);
    else tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace qi

template <int DH, int EMU, bool PAIR>
static int launch_attention_ts(qi::AttParams P, const void* k, const void* v, int batch, int seq_q, int seq_kv, int heads,
                               long long ld_kv, long long kv_bs, long long kv_hs, cudaStream_t stream) {
  using namespace qi;
  using Cfg = TsCfg<DH, PAIR>;
  // K boxes: the key rows of a half that one CTA stages (64, or 32 per CTA of a pair) x 64 head-dim columns;
  // V boxes: 64 keys x 64 head-dim columns
  const uint32_t box_k[4] = {64, 1, (uint32_t)Cfg::K_ROWS, 1};
  const uint32_t box_v[4] = {64, 1, (uint32_t)Cfg::HALF, 1};
  uint64_t dims[4] = {(uint64_t)DH, (uint64_t)heads, (uint64_t)seq_kv, (uint64_t)batch};
  uint64_t str[3] = {(uint64_t)kv_hs * 2, (uint64_t)ld_kv * 2,
                     (uint64_t)(batch > 1 ? kv_bs : ld_kv * (long long)seq_kv) * 2};
  if (int rc = qi_encode_tmap_bf16(&P.tmap_k, k, 4, dims, str, box_k)) return rc;
  if (int rc = qi_encode_tmap_bf16(&P.tmap_v, v, 4, dims, str, box_v)) return rc;
  if (int rc = qi_set_max_smem(reinterpret_cast<const void*>(attention_ts_kernel<DH, EMU, PAIR>), Cfg::SMEM_BYTES))
    return rc;
  const int tiles = (seq_q + ATT_BM - 1) / ATT_BM;
  dim3 grid((tiles + Cfg::NCTA - 1) / Cfg::NCTA * Cfg::NCTA, heads, batch);
  if (qi_launch(attention_ts_kernel<DH, EMU, PAIR>, grid, dim3(ATT_THREADS), Cfg::SMEM_BYTES, stream, P) != cudaSuccess)
    return QI_ERR_CUDA;
  return QI_OK;
}
