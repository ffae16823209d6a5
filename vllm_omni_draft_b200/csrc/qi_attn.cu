// K4: joint text+image flash attention forward for sm_100a (non-causal, no mask, dropout 0).
//
// One CTA owns two 128-query tiles of one (batch, head) and streams 128-key K/V tiles through a
// 2-stage TMA ring.  tcgen05.mma computes S_t = Q_t K^T into TMEM (fp32, 128 columns per tile);
// a softmax warpgroup per Q tile (one thread per query row) reads S_t with tcgen05.ld, does the
// online-softmax update in the exp2 domain, and writes P_t as packed bf16 back over S_t's columns
// (tcgen05.st); tcgen05.mma (A from TMEM) then accumulates O_t += P_t V into TMEM.  The two Q tiles
// ping-pong so the tensor pipe works on one tile while the other tile's softmax runs.
// O is rescaled lazily: only when a row's running max grows by more than 2^8.
//
// Softmax inner loop (the kernel is bound by the S -> P chain, not by the tensor pipe):
//   * row max over 8 independent 3-input-max chains (a single chain is 64 dependent FMNMX3);
//   * x = s*scale - m and the row sum use packed fma.rn.f32x2 / add.rn.f32x2 (half the FMA-pipe slots);
//   * EMU of every 8 element pairs evaluate 2^x on the FMA pipe (round-to-nearest range reduction +
//     degree-3 minimax polynomial, 7.5e-5 relative error, P is rounded to bf16 afterwards) instead of
//     MUFU.EX2, whose 16/clk/SM rate otherwise bounds the softmax at the tensor pipe's own tile time;
//   * SPLIT: P is published in two halves so the first four P*V MMAs overlap the second half's exps;
//   * only the last (partial) K/V tile runs the masked variant of the loop body.
//
// Warps: 0-3 softmax tile 0, 4-7 softmax tile 1, 8 TMA producer, 9 MMA issuer / TMEM owner.
#include "qi_common.cuh"
#include "../../include/qi_b200.h"
#include <cstdlib>
#include <type_traits>

#ifndef QI_ATTN_DEFAULT_VARIANT
#define QI_ATTN_DEFAULT_VARIANT 11   // EMU = 3 of 8 pairs on the FMA pipe, split P
#endif

namespace qi {

constexpr int ATT_BM = 128;   // query rows per tile
constexpr int ATT_BN = 128;   // keys per tile
constexpr int ATT_THREADS = 384;   // 3 warpgroups: softmax 0, softmax 1, {TMA, MMA, 2 idle}
constexpr int ATT_THREADS2 = 640;  // TWO: 2 softmax warpgroups per Q tile (64 keys of every row each) + the same 4 warps
constexpr int ATT_KV_STAGES = 2;

template <int DH>
struct AttCfg {
  static constexpr int SUB = ATT_BM * 128;             // one [128 rows x 64 bf16] swizzled sub-tile
  static constexpr int NSUB = DH / 64;                 // sub-tiles per operand tile
  static constexpr int TILE_BYTES = SUB * NSUB;        // Q, K or V tile
  static constexpr int Q_OFF = 0;
  static constexpr int K_OFF = 2 * TILE_BYTES;
  static constexpr int V_OFF = K_OFF + ATT_KV_STAGES * TILE_BYTES;
  static constexpr int BAR_OFF = V_OFF + ATT_KV_STAGES * TILE_BYTES;
  static constexpr int XCH_OFF = BAR_OFF + 256;        // TWO: float [2 parities][2 tiles][2 halves][128 rows]
  static constexpr int SMEM_BYTES = XCH_OFF + 4096 + 1024;
  // TMEM columns: S0 [0,128) S1 [128,256) O0 [256,256+DH) O1 [384,384+DH)
  static constexpr int S_COL0 = 0, S_COL1 = 128, O_COL0 = 256, O_COL1 = 384;
};

struct AttParams {
  CUtensorMap tmap_q, tmap_k, tmap_v;
  __nv_bfloat16* o;
  long long ld_o, o_bs, o_hs;   // token, batch and head strides of the output (elements)
  const __nv_bfloat16* q;       // TS kernel: Q rows are read straight from global memory into TMEM
#ifdef QI_ATTN_EXPERIMENTS
  int ts_diag;                  // tuning aid (env QI_ATTN_TS_DIAG): 1 = softmax only hands the barriers on, 2 = S load + P store without the math, 3 = as 1 and the P*V issuer does not wait for P
#endif
  long long ld_q, q_bs, q_hs;
  int seq_q, seq_kv, heads;
  float scale_log2;
  int split_p;   // softmax publishes P in two 64-key halves (p_half, then p_full)
  // Ulysses scatter epilogue (n_peer > 0): query row r owned by rank p goes to
  // o_peer[p] + o_chunk_off + b*o_bs + (r - o_prefix[p])*ld_o + head*o_hs   (peer-mapped HBM)
  int n_peer;
  int o_prefix[9];
  long long o_chunk_off;
  __nv_bfloat16* o_peer[8];
  PeerSignal wait;     // flags of the exchange that delivered q/k/v (polled by the TMA thread), or slots == null
  PeerSignal notify;   // flags of the O scatter (signalled by the grid's last CTA), or slots == null
  // tuning aid (qi_attention_set_debug, not part of the ABI): cycle counters summed over CTAs —
  // [0] MMA warp total, [1] waiting v_full, [2] p_half, [3] p_full, [4] k_full, [5] q/k first,
  // [8] softmax warp 0 total, [9] waiting s_full, [10] waiting o_done,
  // [11..15] one-thread softmax phases: S load, row max, O rescale, first half of P, second half
  unsigned long long* dbg;
};

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 2^x for a pair on the FMA/ALU pipes: j = round(x), f = x - j in [-0.5, 0.5], degree-3 minimax
// polynomial for 2^f (max relative error 7.5e-5), exponent added as an integer.  x is clamped to
// >= -126 so the exponent field cannot wrap (masked columns, x = -inf, come out as 2^-126 ~ 0).
template <int DEG = 3>
__device__ __forceinline__ float2 exp2_poly2(float2 x) {
  x.x = fmaxf(x.x, -126.f);
  x.y = fmaxf(x.y, -126.f);
  const float2 t = __fadd2_rn(x, make_float2(12582912.f, 12582912.f));          // 1.5 * 2^23 + j
  const float2 jf = __fadd2_rn(t, make_float2(-12582912.f, -12582912.f));
  const float2 f = __ffma2_rn(jf, make_float2(-1.f, -1.f), x);
  float2 p;
  if (DEG == 3) {
    p = __ffma2_rn(make_float2(0.0551716648f, 0.0551716648f), f, make_float2(0.2426111251f, 0.2426111251f));
    p = __ffma2_rn(p, f, make_float2(0.6932609677f, 0.6932609677f));
    p = __ffma2_rn(p, f, make_float2(0.9999280572f, 0.9999280572f));
  } else {   // degree-2 minimax, max relative error 1.73e-3 (bf16 rounding of P: 3.9e-3): one FFMA2 less per pair
    p = __ffma2_rn(make_float2(0.23842697f, 0.23842697f), f, make_float2(0.7034453f, 0.7034453f));
    p = __ffma2_rn(p, f, make_float2(1.00044307f, 1.00044307f));
  }
  float2 r;
  r.x = __int_as_float(__float_as_int(p.x) + (__float_as_int(t.x) << 23));
  r.y = __int_as_float(__float_as_int(p.y) + (__float_as_int(t.y) << 23));
  return r;
}

// One K/V tile of the online softmax for one query row (one thread): S row from TMEM -> running
// max / lazy O rescale -> P (bf16, packed over S's first 64 columns) + row-sum update.
template <int DH, bool MASK, int EMU, bool SPLIT>
__device__ __forceinline__ void softmax_tile(uint32_t s_addr, uint32_t o_addr, int kv_valid, bool first,
                                             float scale_log2, float& m_ref, float& l_sum,
                                             uint64_t* p_half, uint64_t* p_full, int lane,
                                             unsigned long long* pc = nullptr) {
  uint32_t r[ATT_BN / 32][32];
  long long c_prev = pc ? clock64() : 0;
  auto phase = [&](int i) {   // tuning aid: cycles per phase, thread 0 of the CTA only
    if (pc) {
      const long long now = clock64();
      pc[i] += now - c_prev;
      c_prev = now;
    }
  };
  // STAG (EMU bit 5): the second half of S is loaded from TMEM while the row maximum runs over the first half
  // (tcgen05.wait::ld waits for every outstanding load, so the loads are issued in two groups).
  constexpr bool STAG = (EMU & 32) != 0;
  float mxs[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) mxs[q] = -INFINITY;
  auto mask_and_max = [&](int c) {
    if (MASK) {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (c * 32 + i >= kv_valid) r[c][i] = 0xff800000u;   // -inf: contributes 0
    }
#pragma unroll
    for (int i = 0; i < 16; ++i)
      mxs[i & 7] = fmaxf(fmaxf(mxs[i & 7], __uint_as_float(r[c][2 * i])), __uint_as_float(r[c][2 * i + 1]));
  };
  if (STAG) {
    tmem_ld32(s_addr, r[0]);
    tmem_ld32(s_addr + 32, r[1]);
    tmem_ld_wait();
    tmem_ld32(s_addr + 64, r[2]);
    tmem_ld32(s_addr + 96, r[3]);
    mask_and_max(0);
    mask_and_max(1);
    tmem_ld_wait();
    phase(0);
    mask_and_max(2);
    mask_and_max(3);
  } else {
#pragma unroll
    for (int c = 0; c < ATT_BN / 32; ++c) tmem_ld32(s_addr + c * 32, r[c]);
    tmem_ld_wait();
    phase(0);
#pragma unroll
    for (int c = 0; c < ATT_BN / 32; ++c) mask_and_max(c);
  }
  const float mx = fmaxf(fmaxf(fmaxf(mxs[0], mxs[1]), fmaxf(mxs[2], mxs[3])),
                         fmaxf(fmaxf(mxs[4], mxs[5]), fmaxf(mxs[6], mxs[7])));
  const float m_new = fmaxf(m_ref, mx * scale_log2);
  const bool need = (m_new - m_ref) > 8.0f;   // also true on the first tile (m_ref = -inf)
  phase(1);
  if (__any_sync(0xffffffffu, need)) {
    const float f = need ? fast_exp2(m_ref - m_new) : 1.0f;   // exp2(-inf) = 0 on tile 0
    if (need) {
      l_sum *= f;
      m_ref = m_new;
    }
    if (!first) {
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
  // P = exp2(S*scale - m_ref), packed bf16 over the first 64 columns of S
  constexpr int NE = EMU & 15;                 // pairs of every 8 evaluated on the FMA pipe
  constexpr int DEG = (EMU & 16) ? 2 : 3;      // polynomial degree of that path
  const float2 sc2 = make_float2(scale_log2, scale_log2);
  const float2 nm2 = make_float2(-m_ref, -m_ref);
  float2 acc[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) acc[q] = make_float2(0.f, 0.f);
#pragma unroll
  for (int c = 0; c < ATT_BN / 32; ++c) {
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
    tmem_st16(s_addr + c * 16, pk);
    if (SPLIT && c == ATT_BN / 64 - 1) {
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_half);
      phase(3);
    }
  }
  const float2 a01 = __fadd2_rn(acc[0], acc[1]), a23 = __fadd2_rn(acc[2], acc[3]);
  l_sum += (a01.x + a01.y) + (a23.x + a23.y);
  tmem_st_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(p_full);
  phase(4);
}


// Rendezvous of the two warps that share 32 query rows (TWO): named barrier `id`, 64 threads.
__device__ __forceinline__ void pair_sync(int id) {
  asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory");
}

// TWO: the same update with a row split over two threads (same lane of two warps in different warpgroups, same
// TMEM lane quadrant): this thread owns keys [64*half, 64*half+64) of the tile and columns [half*DH/2, +DH/2) of O.
// The row maximum is combined through shared memory (`xch_mine` / `xch_other`, double-buffered by the caller) and a
// 64-thread named barrier; both threads then take identical rescale decisions.  The row sum stays partial per
// thread and is combined once, in the epilogue.  Half 1 writes P over columns that held half 0's S values: the
// barrier orders that store after half 0's tcgen05.ld (which is waited for before the barrier).
template <int DH, bool MASK, int EMU>
__device__ __forceinline__ void softmax_half_tile(uint32_t s_addr, uint32_t o_addr, int half, int kv_valid,
                                                  bool first, float scale_log2, float& m_ref, float& l_sum,
                                                  float* xch_mine, const float* xch_other, int bar_id,
                                                  uint64_t* p_done, int lane, unsigned long long* pc = nullptr) {
  uint32_t r[2][32];
  long long c_prev = pc ? clock64() : 0;
  auto phase = [&](int i) {
    if (pc) {
      const long long now = clock64();
      pc[i] += now - c_prev;
      c_prev = now;
    }
  };
  tmem_ld32(s_addr + half * 64, r[0]);
  tmem_ld32(s_addr + half * 64 + 32, r[1]);
  tmem_ld_wait();
  phase(0);
  if (MASK) {
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (half * 64 + c * 32 + i >= kv_valid) r[c][i] = 0xff800000u;
  }
  float mxs[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) mxs[q] = -INFINITY;
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int i = 0; i < 16; ++i)
      mxs[i & 7] = fmaxf(fmaxf(mxs[i & 7], __uint_as_float(r[c][2 * i])), __uint_as_float(r[c][2 * i + 1]));
  float mx = fmaxf(fmaxf(fmaxf(mxs[0], mxs[1]), fmaxf(mxs[2], mxs[3])),
                   fmaxf(fmaxf(mxs[4], mxs[5]), fmaxf(mxs[6], mxs[7])));
  *xch_mine = mx;
  phase(1);
  pair_sync(bar_id);
  mx = fmaxf(mx, *xch_other);
  const float m_new = fmaxf(m_ref, mx * scale_log2);
  const bool need = (m_new - m_ref) > 8.0f;
  phase(2);   // TWO: [2] = the exchange (barrier + partner's maximum); the rare O rescale is counted with P
  if (__any_sync(0xffffffffu, need)) {
    const float f = need ? fast_exp2(m_ref - m_new) : 1.0f;
    if (need) {
      l_sum *= f;
      m_ref = m_new;
    }
    if (!first) {
#pragma unroll 1
      for (int c = 0; c < DH / 64; ++c) {
        uint32_t o[32];
        tmem_ld32(o_addr + half * (DH / 2) + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * f);
        tmem_st32(o_addr + half * (DH / 2) + c * 32, o);
      }
    }
  }
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
      if (((i % 8) * EMU) % 8 < EMU) {
        p = exp2_poly2(x);
      } else {
        p.x = fast_exp2(x.x);
        p.y = fast_exp2(x.y);
      }
      acc[i & 3] = __fadd2_rn(acc[i & 3], p);
      pk[i] = pack_bf16(p.x, p.y);
    }
    tmem_st16(s_addr + half * 32 + c * 16, pk);
    phase(3 + c);
  }
  const float2 a01 = __fadd2_rn(acc[0], acc[1]), a23 = __fadd2_rn(acc[2], acc[3]);
  l_sum += (a01.x + a01.y) + (a23.x + a23.y);
  tmem_st_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(p_done);
  phase(4);
}

// COOP: the two softmax warpgroups work on the SAME Q tile at the same time — warpgroup g owns keys [64 g, 64 g + 64)
// of every K/V tile and columns [g DH/2, +DH/2) of O, for both Q tiles in turn — so a tile's S -> P latency, which
// sits on the tile's dependency chain, is that of 64 columns instead of 128, and neither warpgroup idles while the
// other tile's MMAs run.  Unlike TWO there is no per-tile exchange on that chain: the common reference maximum of a
// row is updated from the local maxima of the PREVIOUS K/V tile (the partner's arrives through shared memory, ordered
// by the P / S barriers of the tile in between), so both threads of a row take identical rescale decisions without
// meeting.  Only the first K/V tile, whose reference cannot lag, meets at a named barrier.  The lag costs nothing in
// accuracy (P = 2^(x - m_ref) is exact for any reference; the row sum and O carry the same factor) as long as a
// row's scores do not grow by more than 2^96 within one K/V tile beyond everything seen before: such elements are
// clamped for that one tile (logits 66 nats above the running maximum do not occur behind a QK-RMSNorm).
// P of keys [64 g, 64 g + 64) is written over this thread's own S columns [64 g, 64 g + 32).
template <int DH, bool MASK, int EMU>
__device__ __forceinline__ void softmax_coop_tile(uint32_t s_addr, uint32_t o_addr, int half, int kv_valid, int j,
                                                  float scale_log2, float& m_ref, float& l_part, float& mx_prev,
                                                  float* xch_mine, const float* xch_other,
                                                  const float* xch_other_prev, int bar_id, uint64_t* p_done,
                                                  int lane) {
  uint32_t r[2][32];
  tmem_ld32(s_addr + half * 64, r[0]);
  tmem_ld32(s_addr + half * 64 + 32, r[1]);
  tmem_ld_wait();
  if (MASK) {
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (half * 64 + c * 32 + i >= kv_valid) r[c][i] = 0xff800000u;
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
  *xch_mine = mx;   // read by the partner during the NEXT K/V tile (and at the barrier below for tile 0)
  if (j == 0) {
    pair_sync(bar_id);
    m_ref = fmaxf(mx, *xch_other) * scale_log2;   // key 0 is valid, so the row maximum is finite
  } else {
    const float m_new = fmaxf(m_ref, fmaxf(mx_prev, *xch_other_prev) * scale_log2);
    const bool need = (m_new - m_ref) > 8.0f;
    if (__any_sync(0xffffffffu, need)) {
      const float f = need ? fast_exp2(m_ref - m_new) : 1.0f;
      if (need) {
        l_part *= f;
        m_ref = m_new;
      }
#pragma unroll 1
      for (int c = 0; c < DH / 64; ++c) {
        uint32_t o[32];
        tmem_ld32(o_addr + half * (DH / 2) + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * f);
        tmem_st32(o_addr + half * (DH / 2) + c * 32, o);
      }
    }
    if (__any_sync(0xffffffffu, mx * scale_log2 - m_ref > 96.0f)) {   // see above: keeps 2^x finite
      const float cap = (m_ref + 96.0f) / scale_log2;
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int i = 0; i < 32; ++i) r[c][i] = __float_as_uint(fminf(__uint_as_float(r[c][i]), cap));
    }
  }
  mx_prev = mx;
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
    tmem_st16(s_addr + half * 64 + c * 16, pk);
  }
  const float2 a01 = __fadd2_rn(acc[0], acc[1]), a23 = __fadd2_rn(acc[2], acc[3]);
  l_part += (a01.x + a01.y) + (a23.x + a23.y);
  tmem_st_wait();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(p_done);
}

template <int DH, int EMU, bool SPLIT, bool TWO, bool COOP = false>
__global__ void __launch_bounds__(TWO ? ATT_THREADS2 : ATT_THREADS, 1)
attention_fwd_kernel(const __grid_constant__ AttParams P) {
  using Cfg = AttCfg<DH>;
  static_assert(!TWO || SPLIT, "two threads per row publish P in two halves by construction");
  static_assert(!COOP || SPLIT, "cooperating warpgroups publish the two key halves of P separately");
  constexpr int W_TMA = TWO ? 16 : 8;   // first warp of the last warpgroup
  constexpr int W_MMA = W_TMA + 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::BAR_OFF);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* k_empty = bars + 3;   // [2]
  uint64_t* v_full = bars + 5;    // [2]
  uint64_t* v_empty = bars + 7;   // [2]
  uint64_t* s_full = bars + 9;    // [2]  per Q tile
  uint64_t* p_full = bars + 11;   // [2]
  uint64_t* o_done = bars + 13;   // [2]
  uint64_t* p_half = bars + 15;   // [2]  first 64 keys of P written (SPLIT only)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * (2 * ATT_BM);
  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int n_kv = (P.seq_kv + ATT_BN - 1) / ATT_BN;
  pdl_launch_dependents();

  if (warp == W_TMA && lane == 0) {
    tma_prefetch_desc(&P.tmap_q);
    tma_prefetch_desc(&P.tmap_k);
    tma_prefetch_desc(&P.tmap_v);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4);
      mbar_init(&p_half[i], 4);
      mbar_init(&o_done[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == W_MMA) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  // Register re-balance (setmaxnreg is per warpgroup and must sit inside each role's own branch so
  // that ptxas allocates every role against its own budget): warps 8-11 shrink, softmax warps grow.
  if (warp == W_TMA) {
    // ------------------------------------------------------------------ TMA producer
    if (TWO) asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
    else asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    peer_wait_all_warp(P.wait, lane);
    if (elect_one()) {
      mbar_expect_tx(q_full, 2 * Cfg::TILE_BYTES);
      for (int t = 0; t < 2; ++t)
        for (int s = 0; s < Cfg::NSUB; ++s)
          tma_load_4d(smem + Cfg::Q_OFF + t * Cfg::TILE_BYTES + s * Cfg::SUB, &P.tmap_q, q_full,
                      s * 64, head, q0 + t * ATT_BM, b);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(&k_empty[st], ph ^ 1);
        mbar_expect_tx(&k_full[st], Cfg::TILE_BYTES);
        for (int s = 0; s < Cfg::NSUB; ++s)
          tma_load_4d(smem + Cfg::K_OFF + st * Cfg::TILE_BYTES + s * Cfg::SUB, &P.tmap_k,
                      &k_full[st], s * 64, head, j * ATT_BN, b);
        mbar_wait(&v_empty[st], ph ^ 1);
        mbar_expect_tx(&v_full[st], Cfg::TILE_BYTES);
        for (int s = 0; s < Cfg::NSUB; ++s)
          tma_load_4d(smem + Cfg::V_OFF + st * Cfg::TILE_BYTES + s * Cfg::SUB, &P.tmap_v,
                      &v_full[st], s * 64, head, j * ATT_BN, b);
      }
    }
    __syncwarp();
  } else if (warp == W_MMA) {
    // ------------------------------------------------------------------ MMA issuer
    if (TWO) asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
    else asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    if (elect_one()) {   // elect.sync: ptxas knows a single thread issues (no per-instruction elect loops)
      constexpr uint32_t idesc_qk = make_idesc_bf16(ATT_BM, ATT_BN, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(ATT_BM, DH, 0, 1);   // B (=V) is MN-major
      const uint32_t s_col[2] = {tmem_base + Cfg::S_COL0, tmem_base + Cfg::S_COL1};
      const uint32_t o_col[2] = {tmem_base + Cfg::O_COL0, tmem_base + Cfg::O_COL1};
      const uint32_t q_addr = smem_u32(smem + Cfg::Q_OFF);
      const uint32_t k_addr = smem_u32(smem + Cfg::K_OFF);
      const uint32_t v_addr = smem_u32(smem + Cfg::V_OFF);

      auto issue_qk = [&](int t, int st) {
#pragma unroll
        for (int s = 0; s < Cfg::NSUB; ++s) {
          const uint64_t da = make_smem_desc(q_addr + t * Cfg::TILE_BYTES + s * Cfg::SUB, 0, 1024);
          const uint64_t db = make_smem_desc(k_addr + st * Cfg::TILE_BYTES + s * Cfg::SUB, 0, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mma_ss(s_col[t], da + 2 * k, db + 2 * k, idesc_qk, (s | k) != 0 ? 1u : 0u);
        }
      };
      auto issue_pv = [&](int t, int st, bool accumulate, int k0, int k1) {
#pragma unroll
        for (int k = k0; k < k1; ++k) {
          // V sub-tile rows are keys (128 B each): 16 keys per MMA = 2048 B; the second 64 columns
          // of the head dim live one sub-tile (LBO) further.
          const uint64_t db =
              make_smem_desc(v_addr + st * Cfg::TILE_BYTES + k * 2048, Cfg::SUB, 1024);
          // COOP: P of keys 64..127 sits over S columns [64, 96) (each warpgroup writes over its own S columns)
          const int p_off = COOP ? (k < ATT_BN / 32 ? k * 8 : 64 + (k - ATT_BN / 32) * 8) : k * 8;
          mma_ts(o_col[t], s_col[t] + p_off, db, idesc_pv, (accumulate || k != 0) ? 1u : 0u);
        }
      };

      // The hand-offs P ready -> P*V issued and P*V issued -> next Q*K^T issued sit on each Q tile's dependency chain,
      // and this single thread executes ~10-15 clk per scalar instruction: with the K/V stage and the Q tile as
      // run-time values, ~30 instructions of descriptor arithmetic (and a stack reload) sat between a barrier opening
      // and the first MMA behind it.  The loop is therefore unrolled over both ring stages and both Q tiles — every
      // descriptor, TMEM column and barrier address is loop invariant and stays in uniform registers — and the cycle
      // counters (qi_attention_set_debug) live in their own instantiation of the loop.
      auto run = [&](auto dbg_tag) {
        constexpr bool DBG = decltype(dbg_tag)::value;
        unsigned long long w_v = 0, w_ph = 0, w_pf = 0, w_k = 0, w_q = 0;
        const long long t_begin = DBG ? clock64() : 0;
        auto wait = [&](uint64_t* bar, uint32_t parity, unsigned long long& acc) {
          if (DBG) {
            const long long c0 = clock64();
            mbar_wait(bar, parity);
            acc += clock64() - c0;
          } else {
            mbar_wait(bar, parity);
          }
        };
        wait(q_full, 0, w_q);
        wait(&k_full[0], 0, w_q);
        tc_fence_after();
        issue_qk(0, 0);
        tc_commit(&s_full[0]);
        issue_qk(1, 0);
        tc_commit(&s_full[1]);
        tc_commit(&k_empty[0]);
        for (int j0 = 0; j0 < n_kv; j0 += 2) {
          const uint32_t ph = (j0 >> 1) & 1;   // parity of the ring barriers for K/V tiles j0 (stage 0) and j0 + 1 (stage 1)
#pragma unroll
          for (int st = 0; st < 2; ++st) {
            const int j = j0 + st;
            if (j >= n_kv) break;
            const bool has_next = (j + 1) < n_kv;
            const int stn = st ^ 1;
            const uint32_t phn = st == 0 ? ph : ph ^ 1;   // ((j + 1) >> 1) & 1
            wait(&v_full[st], ph, w_v);
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              if (SPLIT) {
                wait(&p_half[t], st, w_ph);   // j & 1 == st
                tc_fence_after();
                issue_pv(t, st, j > 0, 0, ATT_BN / 32);
                wait(&p_full[t], st, w_pf);
                tc_fence_after();
                issue_pv(t, st, j > 0, ATT_BN / 32, ATT_BN / 16);
              } else {
                wait(&p_full[t], st, w_pf);
                tc_fence_after();
                issue_pv(t, st, j > 0, 0, ATT_BN / 16);
              }
              if (t == 1) tc_commit(&v_empty[st]);
              if (has_next) {
                if (t == 0) {
                  wait(&k_full[stn], phn, w_k);
                  tc_fence_after();
                }
                issue_qk(t, stn);
                tc_commit(&s_full[t]);
                if (t == 1) tc_commit(&k_empty[stn]);
              } else {
                tc_commit(&o_done[t]);
              }
            }
          }
        }
        if (DBG) {
          atomicAdd(&P.dbg[0], (unsigned long long)(clock64() - t_begin));
          atomicAdd(&P.dbg[1], w_v);
          atomicAdd(&P.dbg[2], w_ph);
          atomicAdd(&P.dbg[3], w_pf);
          atomicAdd(&P.dbg[4], w_k);
          atomicAdd(&P.dbg[5], w_q);
        }
      };
      if (P.dbg) run(std::true_type{});
      else run(std::false_type{});
    }
    __syncwarp();
  } else if (warp > W_MMA) {
    if (TWO) asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
    else asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
  } else if (COOP && !TWO) {
    // ------------------------------------------------------------------ softmax, warpgroups cooperating on each Q tile
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int half = warp >> 2;              // which 64 keys of every K/V tile, which half of O's columns
    const int quad = warp & 3;               // TMEM lane quadrant (warp % 4)
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const uint32_t s_addr[2] = {tmem_base + lane_off + Cfg::S_COL0, tmem_base + lane_off + Cfg::S_COL1};
    const uint32_t o_addr[2] = {tmem_base + lane_off + Cfg::O_COL0, tmem_base + lane_off + Cfg::O_COL1};
    float* xch = reinterpret_cast<float*>(smem + Cfg::XCH_OFF);   // [2 parities][2 tiles][2 halves][128 rows]
    const int xrow = quad * 32 + lane;
    const int bar_id = 1 + quad;             // the two warps (one per warpgroup) that share this lane quadrant
    float m_ref[2] = {-INFINITY, -INFINITY}, l_part[2] = {0.f, 0.f}, mx_prev[2] = {-INFINITY, -INFINITY};
    const int n_full = P.seq_kv / ATT_BN;
    const bool prof = P.dbg != nullptr && threadIdx.x == 0;
    unsigned long long w_s = 0;
    const long long t_begin = clock64();
    for (int j = 0; j < n_kv; ++j) {
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        float* xm = xch + (((j & 1) * 2 + t) * 2 + half) * 128 + xrow;
        const float* xo = xch + (((j & 1) * 2 + t) * 2 + (half ^ 1)) * 128 + xrow;
        const float* xop = xch + ((((j & 1) ^ 1) * 2 + t) * 2 + (half ^ 1)) * 128 + xrow;
        uint64_t* p_done = half ? &p_full[t] : &p_half[t];
        const long long c0 = prof ? clock64() : 0;
        mbar_wait(&s_full[t], j & 1);
        if (prof) w_s += clock64() - c0;
        tc_fence_after();
        if (j < n_full) {
          softmax_coop_tile<DH, false, EMU>(s_addr[t], o_addr[t], half, ATT_BN, j, P.scale_log2, m_ref[t], l_part[t],
                                            mx_prev[t], xm, xo, xop, bar_id, p_done, lane);
        } else {
          softmax_coop_tile<DH, true, EMU>(s_addr[t], o_addr[t], half, P.seq_kv - j * ATT_BN, j, P.scale_log2,
                                           m_ref[t], l_part[t], mx_prev[t], xm, xo, xop, bar_id, p_done, lane);
        }
      }
    }
    if (prof) {
      atomicAdd(&P.dbg[8], (unsigned long long)(clock64() - t_begin));
      atomicAdd(&P.dbg[9], w_s);
    }
    constexpr int PPR = DH / 8;
    constexpr int RPI = 32 / PPR;
    constexpr int ITS = 32 / RPI / 2;        // each of the two warps stores half of the 32 rows
#pragma unroll 1
    for (int t = 0; t < 2; ++t) {
      // row sum = both halves (a slot no earlier reader still needs: parity of tile n_kv, which nobody wrote)
      float* xm = xch + (((n_kv & 1) * 2 + t) * 2 + half) * 128 + xrow;
      const float* xo = xch + (((n_kv & 1) * 2 + t) * 2 + (half ^ 1)) * 128 + xrow;
      pair_sync(bar_id);                     // the partner has read this slot's previous value (tile n_kv - 2)
      *xm = l_part[t];
      pair_sync(bar_id);
      const float inv = 1.f / (l_part[t] + *xo);
      mbar_wait(&o_done[t], 0);
      tc_fence_after();
      uint8_t* stage = smem + Cfg::Q_OFF + t * Cfg::TILE_BYTES + quad * (32 * DH * 2);
#pragma unroll 1
      for (int c = 0; c < DH / 64; ++c) {
        uint32_t r[32];
        tmem_ld32(o_addr[t] + half * (DH / 2) + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 o;
          o.x = pack_bf16(__uint_as_float(r[g * 8 + 0]) * inv, __uint_as_float(r[g * 8 + 1]) * inv);
          o.y = pack_bf16(__uint_as_float(r[g * 8 + 2]) * inv, __uint_as_float(r[g * 8 + 3]) * inv);
          o.z = pack_bf16(__uint_as_float(r[g * 8 + 4]) * inv, __uint_as_float(r[g * 8 + 5]) * inv);
          o.w = pack_bf16(__uint_as_float(r[g * 8 + 6]) * inv, __uint_as_float(r[g * 8 + 7]) * inv);
          const int piece = half * (DH / 16) + c * 4 + g;
          *reinterpret_cast<uint4*>(stage + lane * (DH * 2) + ((piece ^ (lane & 7)) << 4)) = o;
        }
      }
      pair_sync(bar_id);
      const int row_base = q0 + t * ATT_BM + quad * 32;
#pragma unroll 4
      for (int it = half * ITS; it < (half + 1) * ITS; ++it) {
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
  } else if (TWO) {
    // ------------------------------------------------------------------ softmax, two threads per query row
    asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
    const int t = warp >> 3;                 // Q tile
    const int half = (warp >> 2) & 1;        // which 64 keys of every K/V tile, which half of O's columns
    const int quad = warp & 3;               // TMEM lane quadrant (warp % 4)
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const uint32_t s_addr = tmem_base + lane_off + (t == 0 ? Cfg::S_COL0 : Cfg::S_COL1);
    const uint32_t o_addr = tmem_base + lane_off + (t == 0 ? Cfg::O_COL0 : Cfg::O_COL1);
    float* xch = reinterpret_cast<float*>(smem + Cfg::XCH_OFF);
    const int xrow = quad * 32 + lane;
    const int bar_id = 1 + t * 4 + quad;
    uint64_t* p_done = half ? &p_full[t] : &p_half[t];
    float m_ref = -INFINITY;
    float l_sum = 0.f;                       // this thread's 64 keys only
    float mx_prev = -INFINITY;               // COOP: this thread's maximum over the previous K/V tile
    const int n_full = P.seq_kv / ATT_BN;
    const bool prof = P.dbg != nullptr && threadIdx.x == 0;
    unsigned long long w_s = 0;
    unsigned long long pcs[5] = {0, 0, 0, 0, 0};
    const long long t_begin = clock64();
    for (int j = 0; j < n_kv; ++j) {
      float* xm = xch + (((j & 1) * 2 + t) * 2 + half) * 128 + xrow;
      const float* xo = xch + (((j & 1) * 2 + t) * 2 + (half ^ 1)) * 128 + xrow;
      const long long c0 = prof ? clock64() : 0;
      mbar_wait(&s_full[t], j & 1);
      if (prof) w_s += clock64() - c0;
      tc_fence_after();
      if (COOP) {   // reference maximum from the previous tile's maxima: no meeting on the tile's chain (see softmax_coop_tile)
        const float* xop = xch + ((((j & 1) ^ 1) * 2 + t) * 2 + (half ^ 1)) * 128 + xrow;
        if (j < n_full)
          softmax_coop_tile<DH, false, EMU>(s_addr, o_addr, half, ATT_BN, j, P.scale_log2, m_ref, l_sum, mx_prev, xm,
                                            xo, xop, bar_id, p_done, lane);
        else
          softmax_coop_tile<DH, true, EMU>(s_addr, o_addr, half, P.seq_kv - j * ATT_BN, j, P.scale_log2, m_ref, l_sum,
                                           mx_prev, xm, xo, xop, bar_id, p_done, lane);
      } else if (j < n_full) {
        softmax_half_tile<DH, false, EMU>(s_addr, o_addr, half, ATT_BN, j == 0, P.scale_log2, m_ref, l_sum, xm, xo,
                                          bar_id, p_done, lane, prof ? pcs : nullptr);
      } else {
        softmax_half_tile<DH, true, EMU>(s_addr, o_addr, half, P.seq_kv - j * ATT_BN, j == 0, P.scale_log2, m_ref,
                                         l_sum, xm, xo, bar_id, p_done, lane);
      }
    }
    {   // row sum = both halves
      float* xm = xch + (((n_kv & 1) * 2 + t) * 2 + half) * 128 + xrow;
      const float* xo = xch + (((n_kv & 1) * 2 + t) * 2 + (half ^ 1)) * 128 + xrow;
      if (COOP) pair_sync(bar_id);           // the partner read this slot (tile n_kv - 2's maximum) during its last tile
      *xm = l_sum;
      pair_sync(bar_id);
      l_sum += *xo;
    }
    if (prof) {
      for (int i = 0; i < 5; ++i) atomicAdd(&P.dbg[11 + i], pcs[i]);
      const long long c0 = clock64();
      mbar_wait(&o_done[t], 0);
      atomicAdd(&P.dbg[10], (unsigned long long)(clock64() - c0));
      atomicAdd(&P.dbg[8], (unsigned long long)(clock64() - t_begin));
      atomicAdd(&P.dbg[9], w_s);
    }
    mbar_wait(&o_done[t], 0);
    tc_fence_after();
    const float inv = 1.f / l_sum;
    constexpr int PPR = DH / 8;
    constexpr int RPI = 32 / PPR;
    uint8_t* stage = smem + Cfg::Q_OFF + t * Cfg::TILE_BYTES + quad * (32 * DH * 2);
#pragma unroll 1
    for (int c = 0; c < DH / 64; ++c) {
      uint32_t r[32];
      tmem_ld32(o_addr + half * (DH / 2) + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 o;
        o.x = pack_bf16(__uint_as_float(r[g * 8 + 0]) * inv, __uint_as_float(r[g * 8 + 1]) * inv);
        o.y = pack_bf16(__uint_as_float(r[g * 8 + 2]) * inv, __uint_as_float(r[g * 8 + 3]) * inv);
        o.z = pack_bf16(__uint_as_float(r[g * 8 + 4]) * inv, __uint_as_float(r[g * 8 + 5]) * inv);
        o.w = pack_bf16(__uint_as_float(r[g * 8 + 6]) * inv, __uint_as_float(r[g * 8 + 7]) * inv);
        const int piece = half * (DH / 16) + c * 4 + g;
        *reinterpret_cast<uint4*>(stage + lane * (DH * 2) + ((piece ^ (lane & 7)) << 4)) = o;
      }
    }
    pair_sync(bar_id);
    const int row_base = q0 + t * ATT_BM + quad * 32;
    constexpr int ITS = 32 / RPI / 2;        // each of the two warps stores half of the 32 rows
#pragma unroll 4
    for (int it = half * ITS; it < (half + 1) * ITS; ++it) {
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
  } else {
    // ------------------------------------------------------------------ softmax warpgroups
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int t = warp >> 2;                 // Q tile
    const int quad = warp & 3;               // TMEM lane quadrant
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const uint32_t s_addr = tmem_base + lane_off + (t == 0 ? Cfg::S_COL0 : Cfg::S_COL1);
    const uint32_t o_addr = tmem_base + lane_off + (t == 0 ? Cfg::O_COL0 : Cfg::O_COL1);
    float m_ref = -INFINITY;   // reference max (scaled, log2 domain) used for P and O
    float l_sum = 0.f;
    const int n_full = P.seq_kv / ATT_BN;   // K/V tiles with all 128 keys valid
    const bool prof = P.dbg != nullptr && threadIdx.x == 0;
    unsigned long long w_s = 0;
    unsigned long long pcs[5] = {0, 0, 0, 0, 0};   // S load, row max, O rescale, first half of P, second half
    const long long t_begin = clock64();
    for (int j = 0; j < n_kv; ++j) {
      const long long c0 = prof ? clock64() : 0;
      mbar_wait(&s_full[t], j & 1);
      if (prof) w_s += clock64() - c0;
      tc_fence_after();
      if (j < n_full) {
        softmax_tile<DH, false, EMU, SPLIT>(s_addr, o_addr, ATT_BN, j == 0, P.scale_log2, m_ref, l_sum,
                                            &p_half[t], &p_full[t], lane, prof ? pcs : nullptr);
      } else {
        softmax_tile<DH, true, EMU, SPLIT>(s_addr, o_addr, P.seq_kv - j * ATT_BN, j == 0, P.scale_log2,
                                           m_ref, l_sum, &p_half[t], &p_full[t], lane);
      }
    }
    // final: O / l -> bf16 -> global
    if (prof) {
      for (int i = 0; i < 5; ++i) atomicAdd(&P.dbg[11 + i], pcs[i]);
      const long long c0 = clock64();
      mbar_wait(&o_done[t], 0);
      atomicAdd(&P.dbg[10], (unsigned long long)(clock64() - c0));
      atomicAdd(&P.dbg[8], (unsigned long long)(clock64() - t_begin));
      atomicAdd(&P.dbg[9], w_s);
    }
    mbar_wait(&o_done[t], 0);
    tc_fence_after();
    const float inv = 1.f / l_sum;
    // O / l -> bf16, staged through this tile's (now idle) Q shared memory so that global stores are
    // coalesced: DH/8 lanes write one row's contiguous DH*2 bytes.  One 16-byte store per lane and row
    // (a lane owns a row in TMEM) turns into one NVLink packet per lane when the row lives on a peer.
    constexpr int PPR = DH / 8;          // 16-byte pieces per row
    constexpr int RPI = 32 / PPR;        // rows per store instruction
    uint8_t* stage = smem + Cfg::Q_OFF + t * Cfg::TILE_BYTES + quad * (32 * DH * 2);
#pragma unroll 1
    for (int c = 0; c < DH / 32; ++c) {
      uint32_t r[32];
      tmem_ld32(o_addr + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 o;
        o.x = pack_bf16(__uint_as_float(r[g * 8 + 0]) * inv, __uint_as_float(r[g * 8 + 1]) * inv);
        o.y = pack_bf16(__uint_as_float(r[g * 8 + 2]) * inv, __uint_as_float(r[g * 8 + 3]) * inv);
        o.z = pack_bf16(__uint_as_float(r[g * 8 + 4]) * inv, __uint_as_float(r[g * 8 + 5]) * inv);
        o.w = pack_bf16(__uint_as_float(r[g * 8 + 6]) * inv, __uint_as_float(r[g * 8 + 7]) * inv);
        *reinterpret_cast<uint4*>(stage + lane * (DH * 2) + (((c * 4 + g) ^ (lane & 7)) << 4)) = o;
      }
    }
    __syncwarp();
    const int row_base = q0 + t * ATT_BM + quad * 32;
#pragma unroll 4
    for (int it = 0; it < 32 / RPI; ++it) {
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
  __syncthreads();
  if (threadIdx.x == 0) peer_notify_last_cta(P.notify, gridDim.x * gridDim.y * gridDim.z);
  if (warp == W_MMA) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace qi

// The product library ships the two-tile kernel above in four softmax forms (variants 0, 3, 8, 11).  Everything
// else that was built and measured slower (two threads per row, cooperating warpgroups, the Q-in-TMEM half-tile
// kernel of tools/experiments/qi_attn_ts.cuh, other exp2 mixes) is compiled only into the experiments library
// (tools/experiments/Makefile, -DQI_ATTN_EXPERIMENTS); the measurements are under profiles/ and in DESIGN.md section 4.
#ifdef QI_ATTN_EXPERIMENTS
#include "qi_attn_ts.cuh"
#endif

using namespace qi;

static int g_attn_variant = -1;   // -1: env QI_ATTN_VARIANT or default; else EMU + 8 * SPLIT (include/qi_b200.h)
#ifdef QI_ATTN_EXPERIMENTS
static bool attn_variant_known(int v) { return v >= 0 && v <= 84; }   // the switch in launch_attention decides
#else
static bool attn_variant_known(int v) { return v == 0 || v == 3 || v == 8 || v == 11; }
#endif
extern "C" int qi_attention_set_variant(int v) {
  if (v != -1 && !attn_variant_known(v)) return QI_ERR_ARG;
  g_attn_variant = v;
  return QI_OK;
}
#ifdef QI_ATTN_EXPERIMENTS
static unsigned long long* g_attn_dbg = nullptr;   // device buffer of >= 16 counters, or null
extern "C" int qi_attention_set_debug(unsigned long long* buf) {
  g_attn_dbg = buf;
  return QI_OK;
}
#endif
static int attn_variant() {
  static const int env = [] {
    const char* e = getenv("QI_ATTN_VARIANT");   // tuning aid: EMU (0,2,3,4) + 8 * SPLIT
    return e ? atoi(e) : -1;
  }();
  const int v = g_attn_variant >= 0 ? g_attn_variant : env;
  return v >= 0 ? v : QI_ATTN_DEFAULT_VARIANT;
}

template <int DH, int EMU, bool SPLIT, bool TWO = false, bool COOP = false>
static int launch_attention_v(const AttParams& P, dim3 grid, cudaStream_t stream) {
  if (int rc = qi_set_max_smem(reinterpret_cast<const void*>(attention_fwd_kernel<DH, EMU, SPLIT, TWO, COOP>), AttCfg<DH>::SMEM_BYTES))
    return rc;
  if (qi_launch(attention_fwd_kernel<DH, EMU, SPLIT, TWO, COOP>, grid, dim3(TWO ? ATT_THREADS2 : ATT_THREADS),
                AttCfg<DH>::SMEM_BYTES, stream, P) != cudaSuccess)
    return QI_ERR_CUDA;
  return QI_OK;
}

template <int DH>
static int launch_attention(const void* q, const void* k, const void* v, void* o, int batch,
                            int seq_q, int seq_kv, int heads, long long ld_q, long long ld_kv,
                            long long ld_o, long long q_bs, long long kv_bs, long long o_bs,
                            long long q_hs, long long kv_hs, long long o_hs,
                            float scale, const qi_peer_scatter_t* dst, cudaStream_t stream) {
  AttParams P;
  memset(&P, 0, sizeof(P));
  if (dst) {
    P.n_peer = dst->n_peers;
    for (int p = 0; p < dst->n_peers; ++p) {
      P.o_peer[p] = (__nv_bfloat16*)dst->base[p];
      P.o_prefix[p] = dst->row_prefix[p];
    }
    P.o_prefix[dst->n_peers] = dst->row_prefix[dst->n_peers];
    P.o_chunk_off = (long long)dst->my_index * dst->chunk_stride;
    for (int i = 0; i < 2; ++i) {
      const qi_peer_signal_t* s = i ? dst->notify : dst->wait;
      PeerSignal& d = i ? P.notify : P.wait;
      if (!s) continue;
      if (!s->slots || !s->state || s->n_peers < 1 || s->n_peers > 8 || s->my_index < 0 || s->my_index >= s->n_peers)
        return QI_ERR_ARG;
      d.slots = reinterpret_cast<uint32_t* const*>(s->slots);
      d.state = s->state;
      d.n = s->n_peers;
      d.me = s->my_index;
      d.timeout_ns = (unsigned long long)(s->timeout_ms ? s->timeout_ms : 2000) * 1000000ull;
    }
  }
  const uint32_t box[4] = {64, 1, ATT_BM, 1};
  {
    uint64_t dims[4] = {(uint64_t)DH, (uint64_t)heads, (uint64_t)seq_q, (uint64_t)batch};
    uint64_t str[3] = {(uint64_t)q_hs * 2, (uint64_t)ld_q * 2,
                       (uint64_t)(batch > 1 ? q_bs : ld_q * (long long)seq_q) * 2};
    int rc = qi_encode_tmap_bf16(&P.tmap_q, q, 4, dims, str, box);
    if (rc) return rc;
  }
  {
    uint64_t dims[4] = {(uint64_t)DH, (uint64_t)heads, (uint64_t)seq_kv, (uint64_t)batch};
    uint64_t str[3] = {(uint64_t)kv_hs * 2, (uint64_t)ld_kv * 2,
                       (uint64_t)(batch > 1 ? kv_bs : ld_kv * (long long)seq_kv) * 2};
    int rc = qi_encode_tmap_bf16(&P.tmap_k, k, 4, dims, str, box);
    if (rc) return rc;
    rc = qi_encode_tmap_bf16(&P.tmap_v, v, 4, dims, str, box);
    if (rc) return rc;
  }
  P.o = (__nv_bfloat16*)o;
  P.q = (const __nv_bfloat16*)q;
  P.ld_q = ld_q;
  P.q_bs = q_bs;
  P.q_hs = q_hs;
  P.ld_o = ld_o;
  P.o_bs = o_bs;
  P.o_hs = o_hs;
  P.seq_q = seq_q;
  P.seq_kv = seq_kv;
  P.heads = heads;
  P.scale_log2 = scale * 1.4426950408889634f;
#ifdef QI_ATTN_EXPERIMENTS
  P.dbg = g_attn_dbg;
  {
    const char* e = getenv("QI_ATTN_TS_DIAG");
    P.ts_diag = e ? atoi(e) : 0;
  }
#else
  P.dbg = nullptr;   // the cycle counters are an experiments-library aid
#endif
  int variant = attn_variant();
#ifdef QI_ATTN_EXPERIMENTS
  // The Q-in-TMEM kernel does not poll folded peer flags (only its K loader would, leaving Q and V unordered:
  // ADVICE round 2); an exchange that needs them runs the shipped kernel.
  if (variant >= 40 && variant < 64 && P.wait.slots) variant = 11;
#endif
  P.split_p = (variant >> 3) & 1;
  dim3 grid((seq_q + 2 * ATT_BM - 1) / (2 * ATT_BM), heads, batch);
  switch (variant) {
    case 0: return launch_attention_v<DH, 0, false>(P, grid, stream);
#ifdef QI_ATTN_EXPERIMENTS
    case 2: return launch_attention_v<DH, 2, false>(P, grid, stream);
#endif
    case 3: return launch_attention_v<DH, 3, false>(P, grid, stream);
#ifdef QI_ATTN_EXPERIMENTS
    case 4: return launch_attention_v<DH, 4, false>(P, grid, stream);
#endif
    case 8: return launch_attention_v<DH, 0, true>(P, grid, stream);
#ifdef QI_ATTN_EXPERIMENTS
    case 9: return launch_attention_v<DH, 1, true>(P, grid, stream);
    case 10: return launch_attention_v<DH, 2, true>(P, grid, stream);
#endif
    case 11: return launch_attention_v<DH, 3, true>(P, grid, stream);
#ifdef QI_ATTN_EXPERIMENTS
    case 12: return launch_attention_v<DH, 4, true>(P, grid, stream);
    case 16: return launch_attention_v<DH, 0, true, true>(P, grid, stream);
    case 18: return launch_attention_v<DH, 2, true, true>(P, grid, stream);
    case 19: return launch_attention_v<DH, 3, true, true>(P, grid, stream);
    case 20: return launch_attention_v<DH, 4, true, true>(P, grid, stream);
    case 24: return launch_attention_v<DH, 16 + 3, true>(P, grid, stream);   // degree-2 polynomial, 3/4/5 of 8 pairs
    case 25: return launch_attention_v<DH, 16 + 4, true>(P, grid, stream);
    case 26: return launch_attention_v<DH, 16 + 5, true>(P, grid, stream);
    case 27: return launch_attention_v<DH, 32 + 3, true>(P, grid, stream);   // 11 with the S load staggered under the row maximum
    case 64: return launch_attention_v<DH, 0, true, false, true>(P, grid, stream);   // cooperating warpgroups
    case 66: return launch_attention_v<DH, 2, true, false, true>(P, grid, stream);
    case 67: return launch_attention_v<DH, 3, true, false, true>(P, grid, stream);
    case 68: return launch_attention_v<DH, 4, true, false, true>(P, grid, stream);
    case 80: return launch_attention_v<DH, 0, true, true, true>(P, grid, stream);    // four warpgroups, no per-tile exchange
    case 82: return launch_attention_v<DH, 2, true, true, true>(P, grid, stream);
    case 83: return launch_attention_v<DH, 3, true, true, true>(P, grid, stream);
    case 84: return launch_attention_v<DH, 4, true, true, true>(P, grid, stream);
#endif
#ifdef QI_ATTN_EXPERIMENTS
#define QI_TS_CASE(V, EMU, PAIR) \
  case V: return launch_attention_ts<DH, EMU, PAIR && DH == 128>(P, k, v, batch, seq_q, seq_kv, heads, ld_kv, kv_bs, kv_hs, stream);
    QI_TS_CASE(40, 0, false) QI_TS_CASE(41, 1, false) QI_TS_CASE(42, 2, false) QI_TS_CASE(43, 3, false) QI_TS_CASE(44, 4, false)
    QI_TS_CASE(49, 16 + 1, false) QI_TS_CASE(50, 16 + 2, false) QI_TS_CASE(51, 16 + 3, false) QI_TS_CASE(52, 16 + 4, false)
    QI_TS_CASE(56, 0, true) QI_TS_CASE(58, 2, true) QI_TS_CASE(59, 3, true) QI_TS_CASE(60, 4, true)
#undef QI_TS_CASE
#endif
    default: return QI_ERR_ARG;
  }
}

static int attention_entry(const void* q, const void* k, const void* v, void* o, int batch,
                           int seq_q, int seq_kv, int heads, int head_dim, long long ld_q,
                           long long ld_kv, long long ld_o, long long q_batch_stride,
                           long long kv_batch_stride, long long o_batch_stride, long long q_hs,
                           long long kv_hs, long long o_hs, float softmax_scale,
                           const qi_peer_scatter_t* dst, cudaStream_t stream) {
  if (!q || !k || !v || (!o && !dst)) return QI_ERR_ARG;
  if (q_hs <= 0 || kv_hs <= 0 || o_hs <= 0 || q_hs % 8 || kv_hs % 8 || o_hs % 8) return QI_ERR_ARG;
  if (batch <= 0 || seq_q <= 0 || seq_kv <= 0 || heads <= 0) return QI_ERR_ARG;
  if (ld_q % 8 || ld_kv % 8 || ld_o % 8 || q_batch_stride % 8 || kv_batch_stride % 8 ||
      o_batch_stride % 8)
    return QI_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) |
       reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(o)) & 15)
    return QI_ERR_ARG;
  if (dst) {
    if (dst->n_peers < 1 || dst->n_peers > 8 || dst->my_index < 0 || dst->my_index >= dst->n_peers ||
        dst->chunk_stride % 8 || dst->row_prefix[0] != 0 || dst->row_prefix[dst->n_peers] != seq_q)
      return QI_ERR_ARG;
    for (int p = 0; p < dst->n_peers; ++p)
      if (!dst->base[p] || (reinterpret_cast<uintptr_t>(dst->base[p]) & 15) ||
          dst->row_prefix[p + 1] < dst->row_prefix[p])
        return QI_ERR_ARG;
  }
  if (qi_num_sms() <= 0) return QI_ERR_CUDA;
  if (head_dim == 128)
    return launch_attention<128>(q, k, v, o, batch, seq_q, seq_kv, heads, ld_q, ld_kv, ld_o,
                                 q_batch_stride, kv_batch_stride, o_batch_stride, q_hs, kv_hs, o_hs,
                                 softmax_scale, dst, stream);
  if (head_dim == 64)
    return launch_attention<64>(q, k, v, o, batch, seq_q, seq_kv, heads, ld_q, ld_kv, ld_o,
                                q_batch_stride, kv_batch_stride, o_batch_stride, q_hs, kv_hs, o_hs,
                                softmax_scale, dst, stream);
  return QI_ERR_UNSUPPORTED;
}

extern "C" int qi_attention_fwd(const void* q, const void* k, const void* v, void* o, int batch,
                                int seq_q, int seq_kv, int heads, int head_dim, long long ld_q,
                                long long ld_kv, long long ld_o, long long q_batch_stride,
                                long long kv_batch_stride, long long o_batch_stride,
                                float softmax_scale, cudaStream_t stream) {
  return attention_entry(q, k, v, o, batch, seq_q, seq_kv, heads, head_dim, ld_q, ld_kv, ld_o,
                         q_batch_stride, kv_batch_stride, o_batch_stride, head_dim, head_dim, head_dim,
                         softmax_scale, nullptr, stream);
}

extern "C" int qi_attention_fwd_scatter(const void* q, const void* k, const void* v, int batch,
                                        int seq_q, int seq_kv, int heads, int head_dim,
                                        long long ld_q, long long ld_kv, long long ld_o,
                                        long long q_batch_stride, long long kv_batch_stride,
                                        long long o_batch_stride, long long q_head_stride,
                                        long long kv_head_stride, long long o_head_stride,
                                        float softmax_scale, const qi_peer_scatter_t* dst,
                                        cudaStream_t stream) {
  if (!dst) return QI_ERR_ARG;
  return attention_entry(q, k, v, nullptr, batch, seq_q, seq_kv, heads, head_dim, ld_q, ld_kv, ld_o,
                         q_batch_stride, kv_batch_stride, o_batch_stride, q_head_stride, kv_head_stride,
                         o_head_stride, softmax_scale, dst, stream);
}
