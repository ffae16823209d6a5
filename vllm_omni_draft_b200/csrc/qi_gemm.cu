// K3: persistent, warp-specialised tcgen05 bf16 GEMM for sm_100a with fused epilogues.
//
//   C[b, m, n] = epi( sum_k A[b, m, k] * W[n, k] + bias[n] )
//
// Roles (320 threads): warp 0 = TMA producer, warp 1 = tcgen05.mma issuer (+ TMEM owner),
// warps 2..9 = epilogue, two per TMEM lane quadrant, each taking half of the tile's columns (TMEM -> registers -> bias/GELU/gated-residual -> bf16 -> global).
// Tile 128 x BN x 64, UMMA 128 x BN x 16, operands K-major in 128B-swizzled shared memory filled by
// TMA, fp32 accumulators double-buffered in TMEM (2 x BN columns) so the epilogue of tile i overlaps
// the main loop of tile i+1.  Up to two problems (img + txt stream) share one persistent launch.
//
// Two kernels share the tile scheduler and the epilogue:
//   gemm_bf16_kernel<BN>       one CTA per SM, tcgen05.mma.cta_group::1, tile 128 x BN
//   gemm_pair_bf16_kernel<BN>  one CTA pair (cluster of 2, one TPC) per tile 256 x BN,
//                              tcgen05.mma.cta_group::2 issued by the leader CTA: each CTA stages its
//                              own 128 rows of A and only HALF of the B tile, so shared-memory fill and
//                              operand-read traffic per FLOP drop by a third.
#include "qi_common.cuh"
#include "../../include/qi_b200.h"
#include <cstdlib>

namespace qi {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_EPI_WARPS = 8;   // two warps per TMEM lane quadrant, each takes half of the tile's columns
constexpr int GEMM_THREADS = 64 + 32 * GEMM_EPI_WARPS;
// Per-warp epilogue staging tile: 32 rows x 64 B, or 32 rows x 128 B for the tile widths the QKV epilogue's
// wide stores use (BN % 128 == 0; the other widths would not fit their smem ring beside it).
template <int BN>
#ifndef QI_QKV_WIDE_STORES
#define QI_QKV_WIDE_STORES 1      // build-time A/B aid: 0 keeps 64-byte row segments everywhere
#endif
// (+1 KB behind the 2 KB staging tile: the fp32 epilogue keeps its warp's gate and bias columns there)
struct EpiWarpBytes { static constexpr int value = (QI_QKV_WIDE_STORES && BN % 128 == 0) ? 4096 : 3072; };

struct GemmProb {
  void* c;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;
  const __nv_bfloat16* gate;
  long long ldc, c_bs, ld_res, res_bs, gate_bs;
  int m, batch, n, k;
  int epilogue;
  int c_fp32;        // C and the residual are fp32 (the residual streams), else bf16
  int gate_fp32;     // gate vectors are fp32 (the fp32 modulation table), else bf16
  int dbg;           // profiling aid (qi_gemm_set_debug): 1 skip residual loads, 2 skip result stores, 4 skip bias/gate loads
  int tiles_m;       // per batch item
  int tiles_m_total; // tiles_m * batch
  int tiles_n;
  int tile_begin;    // first global tile index of this problem
  int k_chunk;       // A's K axis is split into chunks of this many elements (== k when unsplit)
  // QI_EPI_QKV_NORM_ROPE: per-head RMSNorm + MSRoPE on the q and k column blocks, v passes through
  const float2* freqs;          // [>= pos0 + m, head_dim/2] (cos, sin)
  const __nv_bfloat16* wq;      // [head_dim]
  const __nv_bfloat16* wk;
  int pos0, heads, head_dim;
  float qk_eps;
  // optional Ulysses scatter of the head vectors (sc_n > 0): see qi_peer_scatter_t
  int sc_n, sc_hs, sc_s_total, sc_row_offset;
  // scatter tile order (sc_tpg > 0): column tiles are visited destination by destination, starting with
  // the peer after this rank and ending with this rank's own heads (sc_tpg = column tiles per (q|k|v, owner))
  int sc_me, sc_tpg;
  __nv_bfloat16* sc_base[8];
};

struct GemmLaunch {
  CUtensorMap tmap_a[2];
  CUtensorMap tmap_b[2];
  GemmProb prob[2];
  int num_problems;
  int total_tiles;
  int group_m;   // tile rasterisation: supertiles of group_m row tiles x all column tiles (0: column-major order)
  CUtensorMap tmap_sc[8];   // sc_tma: rank p's attention input buffer as [batch*3*heads/P, s_total, head_dim] (TMA store)
  int sc_tma;
  PeerSignal notify;   // slots != null: the launch's last CTA signals the peers (problem 0's scatter->notify)
  PeerSignal wait;     // slots != null: the TMA producers poll before their first load (problem 0's wait)
};

// Host: qi_peer_signal_t -> the by-value device form.
static int fill_signal(PeerSignal& d, const qi_peer_signal_t* s) {
  if (!s) return QI_OK;
  if (!s->slots || !s->state || s->n_peers < 1 || s->n_peers > 8 || s->my_index < 0 || s->my_index >= s->n_peers)
    return QI_ERR_ARG;
  d.slots = reinterpret_cast<uint32_t* const*>(s->slots);
  d.state = s->state;
  d.n = s->n_peers;
  d.me = s->my_index;
  d.timeout_ns = (unsigned long long)(s->timeout_ms ? s->timeout_ms : 2000) * 1000000ull;
  return QI_OK;
}

template <int BN>
struct GemmCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN >= 224) ? 4 : (BN >= 160 ? 5 : 6);
  static constexpr int ACC_STRIDE = (BN > 128) ? 256 : BN;    // TMEM columns between the two accumulators
  static constexpr int TMEM_COLS = (BN == 128) ? 256 : 512;   // power of two
  // per-warp 32 x 64 B staging tile of the epilogue (coalesced residual loads / result stores)
  static constexpr int EPI_WARP_BYTES = EpiWarpBytes<BN>::value;
  static constexpr int EPI_STAGE_BYTES = GEMM_EPI_WARPS * EPI_WARP_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/ + EPI_STAGE_BYTES;
  static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB shared-memory limit");
};

struct TileCoord {
  int p, b, m0, n0;
};
__device__ __forceinline__ TileCoord decode_tile(const GemmLaunch& L, int t, int BN, int TM = BM) {
  TileCoord tc;
  tc.p = (L.num_problems > 1 && t >= L.prob[1].tile_begin) ? 1 : 0;
  const GemmProb& P = L.prob[tc.p];
  const int lt = t - P.tile_begin;
  // L2-aware rasterisation.  The persistent workers take tiles in linear order, so ~`workers` consecutive tiles run
  // together.  Column-major order makes such a wave span every row tile and 2-4 column tiles: the whole A panel
  // is streamed from HBM once per wave (MLP-down at batch 2: 226 MB x 5.8 waves against 376 MB of operands).
  // Supertiles of group_m row tiles x all column tiles make a wave a ~square group_m x workers/group_m patch of
  // the tile grid: it touches group_m A panels and workers/group_m W panels, and successive waves of a supertile
  // reuse its A rows from L2.
  int mt, nt;
  if (L.group_m > 0) {
    const int gsz = L.group_m * P.tiles_n;
    const int g = lt / gsz;
    const int m_base = g * L.group_m;
    const int gm = min(L.group_m, P.tiles_m_total - m_base);
    const int r = lt - g * gsz;
    nt = r / gm;
    mt = m_base + (r - nt * gm);
  } else {
    mt = lt % P.tiles_m_total;
    nt = lt / P.tiles_m_total;
  }
  if (P.sc_tpg > 0) {
    // Ulysses scatter: rank r visits owners r+1, r+2, ..., r-1 (every rank sends to a different peer at any
    // moment, so no GPU's NVLink ingress is a hot spot), q, k and v of each; its own heads come last, so the
    // final wave's epilogue stores are local while the NVLink writes of the earlier waves drain.
    const int lg = nt / P.sc_tpg, w = nt - lg * P.sc_tpg;
    const int R = P.sc_n - 1;
    int j3, owner;
    if (lg < 3 * R) {
      const int k = lg / 3;
      j3 = lg - 3 * k;
      owner = P.sc_me + 1 + k;
      if (owner >= P.sc_n) owner -= P.sc_n;
    } else {
      j3 = lg - 3 * R;
      owner = P.sc_me;
    }
    nt = (j3 * P.sc_n + owner) * P.sc_tpg + w;
  }
  tc.b = mt / P.tiles_m;
  tc.m0 = (mt % P.tiles_m) * TM;
  tc.n0 = nt * BN;
  return tc;
}

__device__ __forceinline__ float gelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * x * (1.f + t);
}

// One epilogue warp's share of a tile: 32 accumulator rows (one per lane) x BN columns, read from
// TMEM 32 columns at a time; `release` runs right after the last TMEM read so the MMA warp can reuse
// the accumulator while the math and the stores of the last chunk are still in flight.
// A lane owns a ROW of the accumulator, so direct global accesses would be one 16-byte access per
// lane and row (32 different lines per instruction).  Every 32 x 64 B chunk is therefore staged
// through a per-warp shared-memory tile (XOR-swizzled, conflict-free both ways): residual rows are
// loaded and result rows stored with 4 lanes per row (64 contiguous bytes, 8 rows per instruction),
// and the residual of chunk c+1 is prefetched into registers while chunk c is computed.
template <int BN, typename Release>
__device__ __forceinline__ void epilogue_rows(const GemmProb& P, int b, int row, int n0,
                                              uint32_t taddr, int half, uint8_t* stage, Release release) {
  const int lane = threadIdx.x & 31;
  const int row0 = row - lane;
  __nv_bfloat16* cbase = reinterpret_cast<__nv_bfloat16*>(P.c) + (size_t)b * P.c_bs;
  const __nv_bfloat16* rbase = P.residual ? P.residual + (size_t)b * P.res_bs : nullptr;
  const __nv_bfloat16* grow = P.gate ? P.gate + (size_t)b * P.gate_bs : nullptr;
  const bool gated = P.epilogue == QI_EPI_GATED_RESIDUAL;
  constexpr int NCH = BN / 32, SPLIT = (NCH + 1) / 2;   // 32-column chunks: warp half 0 takes [0, SPLIT)
  const int c_begin = half ? SPLIT : 0;
  const int c_end = half ? NCH : SPLIT;
  const int piece = lane & 3;          // cooperative pattern: row it*8 + lane/4, 16-byte piece lane%4
  const int rsub = lane >> 2;
  uint4 rbuf[4];
  auto load_residual = [&](int c) {
    const int n = n0 + c * 32 + piece * 8;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int rr = row0 + it * 8 + rsub;
      rbuf[it] = (rr < P.m && n < P.n) ? *reinterpret_cast<const uint4*>(rbase + (size_t)rr * P.ld_res + n)
                                       : make_uint4(0, 0, 0, 0);
    }
  };
  if (gated) load_residual(c_begin);
#pragma unroll 1
  for (int c = c_begin; c < c_end; ++c) {
    if (gated) {
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int rr = it * 8 + rsub;
        *reinterpret_cast<uint4*>(stage + rr * 64 + ((piece ^ ((rr >> 1) & 3)) << 4)) = rbuf[it];
      }
      __syncwarp();
      if (c + 1 < c_end) load_residual(c + 1);
    }
    uint32_t r[32];
    tmem_ld32(taddr + c * 32, r);
    tmem_ld_wait();
    if (c == c_end - 1) release();
    const int n_base = n0 + c * 32;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int n = n_base + g * 8;
      uint4* slot = reinterpret_cast<uint4*>(stage + lane * 64 + ((g ^ ((lane >> 1) & 3)) << 4));
      if (n >= P.n) break;
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]);
      if (P.bias) {
        const uint4 bu = *reinterpret_cast<const uint4*>(P.bias + n);
        const float2 b0 = unpack_bf16(bu.x), b1 = unpack_bf16(bu.y), b2 = unpack_bf16(bu.z),
                     b3 = unpack_bf16(bu.w);
        v[0] += b0.x; v[1] += b0.y; v[2] += b1.x; v[3] += b1.y;
        v[4] += b2.x; v[5] += b2.y; v[6] += b3.x; v[7] += b3.y;
      }
      if (P.epilogue == QI_EPI_BIAS_GELU_TANH) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = gelu_tanh(v[j]);
      } else if (gated) {
        float gt[8];
        if (P.gate_fp32) {
          const float4* gp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(P.gate) + (size_t)b * P.gate_bs + n);
          const float4 ga = gp[0], gb = gp[1];
          gt[0] = ga.x; gt[1] = ga.y; gt[2] = ga.z; gt[3] = ga.w; gt[4] = gb.x; gt[5] = gb.y; gt[6] = gb.z; gt[7] = gb.w;
        } else {
          const uint4 gu = *reinterpret_cast<const uint4*>(grow + n);
          const float2 g0 = unpack_bf16(gu.x), g1 = unpack_bf16(gu.y), g2 = unpack_bf16(gu.z),
                       g3 = unpack_bf16(gu.w);
          gt[0] = g0.x; gt[1] = g0.y; gt[2] = g1.x; gt[3] = g1.y; gt[4] = g2.x; gt[5] = g2.y; gt[6] = g3.x; gt[7] = g3.y;
        }
        const uint4 ru = *slot;          // this row's residual piece (zero outside the matrix)
        const float2 r0 = unpack_bf16(ru.x), r1 = unpack_bf16(ru.y), r2 = unpack_bf16(ru.z),
                     r3 = unpack_bf16(ru.w);
        const float rs[8] = {r0.x, r0.y, r1.x, r1.y, r2.x, r2.y, r3.x, r3.y};
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = rs[j] + gt[j] * v[j];
      }
      *slot = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                         pack_bf16(v[6], v[7]));
    }
    __syncwarp();
    {
      const int n = n_base + piece * 8;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int rr = it * 8 + rsub;
        const uint4 o = *reinterpret_cast<const uint4*>(stage + rr * 64 + ((piece ^ ((rr >> 1) & 3)) << 4));
        if (row0 + rr < P.m && n < P.n)
          *reinterpret_cast<uint4*>(cbase + (size_t)(row0 + rr) * P.ldc + n) = o;
      }
    }
    __syncwarp();
  }
}

// fp32 variant of epilogue_rows (P.c_fp32): C and the residual R are fp32 — the MMDiT residual streams h / e
// (QI_EPI_BIAS for img_in / txt_in, QI_EPI_GATED_RESIDUAL for the out-projection and MLP-down).  Same staging scheme
// with 16-column sub-chunks, so a staged row is again 64 bytes (16 fp32): residual rows are loaded and result rows
// stored as 64 contiguous bytes per 4 lanes, 8 rows per instruction.
// The residual does not depend on the accumulator, and its loads are what the epilogue waits for (one sub-chunk
// in flight per warp bounds an SM at ~16 KB / latency): the first PF sub-chunks are requested BEFORE the warp
// waits for the accumulator (`wait`), and PF sub-chunks stay in flight after that.  With one sub-chunk in flight
// the out-projection (K = 3072: 13 us of MMAs per 256 x 256 tile) ran at 941 TFLOP/s against 1 374 with bf16
// streams — the epilogue, not the tensor pipe, set the tile time.
template <int BN, typename Wait, typename Release>
__device__ __forceinline__ void epilogue_rows_f32(const GemmProb& P, int b, int row, int n0,
                                                  uint32_t taddr, int half, uint8_t* stage, Wait wait, Release release) {
  const int lane = threadIdx.x & 31;
  const int row0 = row - lane;
  float* cbase = reinterpret_cast<float*>(P.c) + (size_t)b * P.c_bs;
  const float* rbase = P.residual ? reinterpret_cast<const float*>(P.residual) + (size_t)b * P.res_bs : nullptr;
  const __nv_bfloat16* grow = P.gate ? P.gate + (size_t)b * P.gate_bs : nullptr;
  const bool gated = P.epilogue == QI_EPI_GATED_RESIDUAL;
  constexpr int NCH = BN / 16, SPLIT = 2 * ((BN / 32 + 1) / 2);   // same column split between the two warps as bf16
  constexpr int MAXI = SPLIT > NCH - SPLIT ? SPLIT : NCH - SPLIT;
  constexpr int PF = 3;                                            // residual sub-chunks in flight per warp
  const int c_begin = half ? SPLIT : 0;
  const int cnt = half ? NCH - SPLIT : SPLIT;
  const int piece = lane & 3;          // cooperative pattern: row it*8 + lane/4, 16-byte piece (4 fp32) lane%4
  const int rsub = lane >> 2;
  float4 rbuf[PF][4];
  auto load_residual = [&](int c, float4 (&dst)[4]) {
    const int n = n0 + c * 16 + piece * 4;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int rr = row0 + it * 8 + rsub;
      dst[it] = (rr < P.m && n < P.n && !(P.dbg & 1)) ? *reinterpret_cast<const float4*>(rbase + (size_t)rr * P.ld_res + n)
                                                      : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  if (gated) {
#pragma unroll
    for (int p = 0; p < PF; ++p)
      if (p < cnt) load_residual(c_begin + p, rbuf[p]);
  }
  // This warp's gate and bias columns (<= 128 of each) go to shared memory once per tile; every lane then reads
  // them as broadcasts.  Loaded per sub-chunk from global they were the single largest cost of this epilogue
  // (out-projection 211 -> 160 us with them removed): the 2 KB-per-sub-chunk residual stream evicts them from the
  // ~28 KB of L1 that the 200 KB operand ring leaves, so every sub-chunk waited on an L2 round trip.
  float* gsm = reinterpret_cast<float*>(stage + 2048);
  float* bsm = gsm + 128;
  {
    const int j = lane * 4;
    const int n = n0 + c_begin * 16 + j;
    if (j < cnt * 16) {
      float4 gt = make_float4(0.f, 0.f, 0.f, 0.f), bb = gt;
      if (n < P.n && !(P.dbg & 4)) {
        if (gated) {
          if (P.gate_fp32) {
            gt = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(P.gate) + (size_t)b * P.gate_bs + n);
          } else {
            const uint2 gu = *reinterpret_cast<const uint2*>(grow + n);
            const float2 g0 = unpack_bf16(gu.x), g1 = unpack_bf16(gu.y);
            gt = make_float4(g0.x, g0.y, g1.x, g1.y);
          }
        }
        if (P.bias) {
          const uint2 bu = *reinterpret_cast<const uint2*>(P.bias + n);
          const float2 b0 = unpack_bf16(bu.x), b1 = unpack_bf16(bu.y);
          bb = make_float4(b0.x, b0.y, b1.x, b1.y);
        }
      }
      *reinterpret_cast<float4*>(gsm + j) = gt;
      *reinterpret_cast<float4*>(bsm + j) = bb;
    }
  }
  __syncwarp();
  wait();
#pragma unroll
  for (int i = 0; i < MAXI; ++i) {
    if (i >= cnt) break;
    const int c = c_begin + i;
    if (gated) {
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int rr = it * 8 + rsub;
        *reinterpret_cast<float4*>(stage + rr * 64 + ((piece ^ ((rr >> 1) & 3)) << 4)) = rbuf[i % PF][it];
      }
      __syncwarp();
      if (i + PF < cnt) load_residual(c + PF, rbuf[i % PF]);
    }
    uint32_t r[16];
    tmem_ld16(taddr + c * 16, r);
    tmem_ld_wait();
    if (i == cnt - 1) release();
    const int n_base = n0 + c * 16;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int n = n_base + g * 4;
      float4* slot = reinterpret_cast<float4*>(stage + lane * 64 + ((g ^ ((lane >> 1) & 3)) << 4));
      if (n >= P.n) break;
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = __uint_as_float(r[g * 4 + j]);
      const float4 bb = *reinterpret_cast<const float4*>(bsm + i * 16 + g * 4);
      v[0] += bb.x; v[1] += bb.y; v[2] += bb.z; v[3] += bb.w;
      if (gated) {
        const float4 gt = *reinterpret_cast<const float4*>(gsm + i * 16 + g * 4);
        const float4 rs = *slot;          // this row's residual piece (zero outside the matrix)
        v[0] = fmaf(gt.x, v[0], rs.x); v[1] = fmaf(gt.y, v[1], rs.y);
        v[2] = fmaf(gt.z, v[2], rs.z); v[3] = fmaf(gt.w, v[3], rs.w);
      }
      *slot = make_float4(v[0], v[1], v[2], v[3]);
    }
    __syncwarp();
    {
      const int n = n_base + piece * 4;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int rr = it * 8 + rsub;
        const float4 o = *reinterpret_cast<const float4*>(stage + rr * 64 + ((piece ^ ((rr >> 1) & 3)) << 4));
        if (row0 + rr < P.m && n < P.n && !(P.dbg & 2))
          *reinterpret_cast<float4*>(cbase + (size_t)(row0 + rr) * P.ldc + n) = o;
      }
    }
    __syncwarp();
  }
}

// QKV epilogue: the tile's BN columns are BN/DH whole head vectors of q, k or v.  Per row (one thread):
// y = bf16(acc + bias); for q/k heads: RMSNorm over the DH columns (two passes over TMEM: sum of
// squares, then normalise), * norm weight, MSRoPE rotation of adjacent (even, odd) pairs with the row's
// (cos, sin) table -- rounding points as in qknorm_rope_kernel (upstream's eager order).  The head
// vector is stored in place into C, or (sc_n > 0) straight into the head owner's attention input buffer
// over NVLink: the head<->sequence all-to-all happens tile by tile under the GEMM's main loop.
template <int BN, int DH, typename Release>
__device__ __forceinline__ void epilogue_rows_qkv(const GemmProb& P, int b, int row, int n0,
                                                  uint32_t taddr, int half, uint8_t* stage, Release release,
                                                  const CUtensorMap* sc_maps) {
  constexpr int HPT = BN / DH;   // head vectors per tile row
  constexpr bool WIDE = EpiWarpBytes<BN>::value >= 4096;   // 128-byte row segments (else 64-byte)
  const bool row_ok = row < P.m;
  const int lane = threadIdx.x & 31;
  const int row0 = row - lane;   // first row of this warp's 32-row slab
  const int d_model = P.heads * DH;
  const float2* fr = P.freqs + (size_t)(P.pos0 + (row_ok ? row : 0)) * (DH / 2);
  bool released = false;
#pragma unroll 1
  for (int hh = half; hh < HPT; hh += 2) {      // the quadrant's two warps alternate head vectors
    const int nh = n0 + hh * DH;
    if (nh >= P.n) break;
    const int j3 = nh / d_model;                 // 0 q, 1 k, 2 v
    const int h = (nh - j3 * d_model) / DH;
    // Destination of row r's head vector: dst_base + r * dst_ld (in place in C, or the head owner's
    // [batch, 3, heads/P, s_total, DH] attention input buffer over NVLink)
    __nv_bfloat16* dst_base;
    size_t dst_ld;
    const CUtensorMap* tmap = nullptr;   // TMA store of whole 32-row slabs (scatter, WIDE staging)
    int tma_slab = 0;
    if (P.sc_n > 0) {
      const int p = h / P.sc_hs;
      const int hl = h - p * P.sc_hs;
      if (WIDE && sc_maps != nullptr && row0 + 32 <= P.m) {
        tmap = sc_maps + p;
        tma_slab = (b * 3 + j3) * P.sc_hs + hl;
      }
      // head-major destination [batch, 3, heads/P, s_total, DH]: one head's rows are adjacent
      dst_ld = (size_t)DH;
      dst_base = P.sc_base[p] +
                 ((((size_t)b * 3 + j3) * P.sc_hs + hl) * P.sc_s_total + P.sc_row_offset + P.pos0) * DH;
    } else {
      dst_ld = (size_t)P.ldc;
      dst_base = reinterpret_cast<__nv_bfloat16*>(P.c) + (size_t)b * P.c_bs + nh;
    }
    const uint32_t th = taddr + hh * DH;
    float rstd = 0.f;
    if (j3 < 2) {
      float ss = 0.f;
#pragma unroll 1
      for (int c = 0; c < DH / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(th + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float bv[8] = {0, 0, 0, 0, 0, 0, 0, 0};
          if (P.bias) {
            const uint4 bu = *reinterpret_cast<const uint4*>(P.bias + nh + c * 32 + g * 8);
            const float2 b0 = unpack_bf16(bu.x), b1 = unpack_bf16(bu.y), b2 = unpack_bf16(bu.z),
                         b3 = unpack_bf16(bu.w);
            bv[0] = b0.x; bv[1] = b0.y; bv[2] = b1.x; bv[3] = b1.y;
            bv[4] = b2.x; bv[5] = b2.y; bv[6] = b3.x; bv[7] = b3.y;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float v = bf16_round(__uint_as_float(r[g * 8 + j]) + bv[j]);
            ss = fmaf(v, v, ss);
          }
        }
      }
      rstd = rsqrtf(ss * (1.f / DH) + P.qk_eps);
    }
    const __nv_bfloat16* wn = (j3 == 0) ? P.wq : P.wk;
#pragma unroll 1
    for (int c = 0; c < DH / 32; ++c) {
      uint32_t r[32];
      tmem_ld32(th + c * 32, r);
      tmem_ld_wait();
      if (hh + 2 >= HPT && c == DH / 32 - 1) {
        release();
        released = true;
      }
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const int col = c * 32 + g * 8;     // column inside the head vector
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]);
        if (P.bias) {
          const uint4 bu = *reinterpret_cast<const uint4*>(P.bias + nh + col);
          const float2 b0 = unpack_bf16(bu.x), b1 = unpack_bf16(bu.y), b2 = unpack_bf16(bu.z),
                       b3 = unpack_bf16(bu.w);
          v[0] += b0.x; v[1] += b0.y; v[2] += b1.x; v[3] += b1.y;
          v[4] += b2.x; v[5] += b2.y; v[6] += b3.x; v[7] += b3.y;
        }
        if (j3 < 2) {
          const uint4 wu = *reinterpret_cast<const uint4*>(wn + col);
          const float2 w0 = unpack_bf16(wu.x), w1 = unpack_bf16(wu.y), w2 = unpack_bf16(wu.z),
                       w3 = unpack_bf16(wu.w);
          const float w[8] = {w0.x, w0.y, w1.x, w1.y, w2.x, w2.y, w3.x, w3.y};
          const float4 f01 = *reinterpret_cast<const float4*>(fr + col / 2);
          const float4 f23 = *reinterpret_cast<const float4*>(fr + col / 2 + 2);
          const float cs[4] = {f01.x, f01.z, f23.x, f23.z};
          const float sn[4] = {f01.y, f01.w, f23.y, f23.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float a = bf16_round(bf16_round(bf16_round(v[2 * j]) * rstd) * w[2 * j]);
            const float bb = bf16_round(bf16_round(bf16_round(v[2 * j + 1]) * rstd) * w[2 * j + 1]);
            v[2 * j] = a * cs[j] - bb * sn[j];
            v[2 * j + 1] = a * sn[j] + bb * cs[j];
          }
        }
        // stage: piece g (16 B) of this lane's row, XOR-swizzled so neither side has bank conflicts
        const uint4 o = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]),
                                   pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
        if constexpr (WIDE) {
          if (sc_maps != nullptr && g == 0 && (c & 1) == 0) {
            // the staging tile may still be the source of the previous bulk store
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
          }
          const int gp = (c & 1) * 4 + g;      // piece inside the 128-byte (two-chunk) row segment
          *reinterpret_cast<uint4*>(stage + lane * 128 + ((gp ^ (lane & 7)) << 4)) = o;
        } else {
          *reinterpret_cast<uint4*>(stage + lane * 64 + ((g ^ ((lane >> 1) & 3)) << 4)) = o;
        }
      }
      if constexpr (WIDE) {
        // Two chunks make one full 128-byte line per row: 8 lanes write a row's line, 4 rows per instruction.
        // Over NVLink the packet count halves against 64-byte segments (the scatter is bound by the peer
        // write rate, not by the tensor pipe, once the head vectors leave the GPU).
        if ((c & 1) && tmap != nullptr) {
          // The staging tile is [32 rows][128 B] with the 16-byte pieces XOR-swizzled by (row & 7): exactly TMA's
          // SWIZZLE_128B.  One bulk tensor store moves the slab's 4 KB to the head owner's HBM (32 rows of one
          // head are adjacent there), instead of 8 st.global.v4 per lane through the LSU and over NVLink.
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) {
            asm volatile(
                "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];\n"
                "cp.async.bulk.commit_group;" ::"l"(reinterpret_cast<uint64_t>(tmap)),
                "r"((c - 1) * 32), "r"(P.sc_row_offset + P.pos0 + row0), "r"(tma_slab), "r"(smem_u32(stage))
                : "memory");
          }
        } else if (c & 1) {
          __syncwarp();
#pragma unroll 4
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + (lane >> 3);
            const int piece = lane & 7;
            const uint4 o = *reinterpret_cast<const uint4*>(stage + rr * 128 + ((piece ^ (rr & 7)) << 4));
            if (row0 + rr < P.m)
              *reinterpret_cast<uint4*>(dst_base + (size_t)(row0 + rr) * dst_ld + (c - 1) * 32 + piece * 8) = o;
          }
          __syncwarp();
        }
      } else {
        __syncwarp();
        // coalesced stores: 4 lanes write one row's 64 contiguous bytes, 8 rows per instruction (peer
        // writes over NVLink need this: one 16-byte store per lane and row is a packet per lane)
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int rr = it * 8 + (lane >> 2);
          const int piece = lane & 3;
          const uint4 o = *reinterpret_cast<const uint4*>(stage + rr * 64 + ((piece ^ ((rr >> 1) & 3)) << 4));
          if (row0 + rr < P.m)
            *reinterpret_cast<uint4*>(dst_base + (size_t)(row0 + rr) * dst_ld + c * 32 + piece * 8) = o;
        }
        __syncwarp();
      }
    }
  }
  if (!released) release();
}

// Epilogue dispatch for one warp's 32 rows of a tile.  `wait` blocks until the tile's accumulator is complete; the
// variants that can start work earlier (fp32 residual prefetch) call it late, the others first thing.
template <int BN, typename Wait, typename Release>
__device__ __forceinline__ void epilogue_tile(const GemmProb& P, int b, int row, int n0, uint32_t taddr,
                                              int half, uint8_t* stage, Wait wait, Release release,
                                              const CUtensorMap* sc_maps = nullptr) {
  if (P.c_fp32 && P.epilogue != QI_EPI_QKV_NORM_ROPE) {
    epilogue_rows_f32<BN>(P, b, row, n0, taddr, half, stage, wait, release);
    return;
  }
  wait();
  if (P.epilogue == QI_EPI_QKV_NORM_ROPE) {
    if constexpr (BN % 128 == 0) {
      if (P.head_dim == 128) {
        epilogue_rows_qkv<BN, 128>(P, b, row, n0, taddr, half, stage, release, sc_maps);
        return;
      }
    }
    if constexpr (BN % 64 == 0)
      epilogue_rows_qkv<BN, 64>(P, b, row, n0, taddr, half, stage, release, nullptr);   // host guarantees head_dim == 64 here
    return;
  }
  epilogue_rows<BN>(P, b, row, n0, taddr, half, stage, release);
}

template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ GemmLaunch L) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::EPI_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + Cfg::STAGES;
  uint64_t* tmem_full = empty_bar + Cfg::STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    for (int p = 0; p < L.num_problems; ++p) {
      tma_prefetch_desc(&L.tmap_a[p]);
      tma_prefetch_desc(&L.tmap_b[p]);
    }
    for (int s = 0; s < Cfg::STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], GEMM_EPI_WARPS);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();   // everything above overlapped the previous kernel's tail

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    peer_wait_all_warp(L.wait, lane);
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < L.total_tiles; t += gridDim.x) {
        const TileCoord tc = decode_tile(L, t, BN);
        const int kblocks = L.prob[tc.p].k / BK;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
          uint8_t* sb = sa + Cfg::A_BYTES;
          mbar_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
          const int kidx = kb * BK;
          const int kc = L.prob[tc.p].k_chunk;
          tma_load_4d(sa, &L.tmap_a[tc.p], &full_bar[stage], kidx % kc, tc.m0, kidx / kc, tc.b);
          tma_load_2d(sb, &L.tmap_b[tc.p], &full_bar[stage], kb * BK, tc.n0);
          if (++stage == Cfg::STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < L.total_tiles; t += gridDim.x) {
        const TileCoord tc = decode_tile(L, t, BN);
        const int kblocks = L.prob[tc.p].k / BK;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * Cfg::ACC_STRIDE;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t sb = sa + Cfg::A_BYTES;
          const uint64_t da = make_smem_desc(sa, 0, 1024);
          const uint64_t db = make_smem_desc(sb, 0, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // advance 16 bf16 = 32 bytes inside the 128B swizzle row: +2 in the (>>4) address field
            mma_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          tc_commit(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
          if (++stage == Cfg::STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        tc_commit(&tmem_full[acc]);  // accumulator complete -> epilogue
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue warps (2..9)
    const int q = warp & 3;  // TMEM lane quadrant this warp may access
    uint8_t* epi_stage = smem + Cfg::STAGES * Cfg::STAGE_BYTES + (warp - 2) * Cfg::EPI_WARP_BYTES;   // 1024-aligned
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < L.total_tiles; t += gridDim.x) {
      const TileCoord tc = decode_tile(L, t, BN);
      const GemmProb& P = L.prob[tc.p];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * Cfg::ACC_STRIDE;
      epilogue_tile<BN>(P, tc.b, tc.m0 + q * 32 + lane, tc.n0, taddr, (warp - 2) >> 2, epi_stage, [&]() {
        mbar_wait(&tmem_full[acc], acc_phase);
        tc_fence_after();
      }, [&]() {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      }, L.sc_tma ? L.tmap_sc : nullptr);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (L.sc_tma && lane == 0) {   // this lane's bulk stores have been written before the CTA signals / exits
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
      asm volatile("fence.proxy.async;" ::: "memory");
    }
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) peer_notify_last_cta(L.notify, gridDim.x);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------
// CTA-pair kernel: cluster of 2 CTAs (one TPC) per 256 x BN tile, tcgen05.mma.cta_group::2.
//   CTA r of the pair stages A rows [m0 + 128 r, +128) and B rows [n0 + BN/2 r, +BN/2) per k-block;
//   both CTAs' TMA transactions complete on the LEADER's full barrier; the leader's MMA thread issues
//   the 256 x BN x 16 MMAs (each CTA's tensor core reads its A rows and both B halves) and commits
//   with a multicast arrive to the empty / accumulator-full barriers of both CTAs; each CTA's epilogue
//   warps drain their own 128 accumulator rows from their own TMEM and release the accumulator on the
//   leader's tmem_empty barrier (count 8 = 4 warps x 2 CTAs).
// ---------------------------------------------------------------------------------------------
template <int BN, int STG = 0>
struct GemmPairCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = (BN / 2) * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = STG ? STG : ((BN >= 224) ? 6 : 7);
  static constexpr int ACC_STRIDE = (BN > 128) ? 256 : BN;
  static constexpr int TMEM_COLS = (BN == 128) ? 256 : 512;
  static constexpr int EPI_WARP_BYTES = EpiWarpBytes<BN>::value;
  static constexpr int EPI_STAGE_BYTES = GEMM_EPI_WARPS * EPI_WARP_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256 + EPI_STAGE_BYTES;
  static_assert(SMEM_BYTES <= 232448, "exceeds the 227 KB shared-memory limit");
};

template <int BN, int STG = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_pair_bf16_kernel(const __grid_constant__ GemmLaunch L) {
  using Cfg = GemmPairCfg<BN, STG>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::EPI_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + Cfg::STAGES;
  uint64_t* tmem_full = empty_bar + Cfg::STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    for (int p = 0; p < L.num_problems; ++p) {
      tma_prefetch_desc(&L.tmap_a[p]);
      tma_prefetch_desc(&L.tmap_b[p]);
    }
    for (int s = 0; s < Cfg::STAGES; ++s) {
      mbar_init(&full_bar[s], 1);    // leader's producer: arrive.expect_tx for both CTAs' bytes
      mbar_init(&empty_bar[s], 1);   // multicast commit from the leader's MMA thread
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);   // multicast commit
      mbar_init(&tmem_empty[a], 2 * GEMM_EPI_WARPS);  // the epilogue warps of both CTAs (used in the leader only)
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair<Cfg::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    peer_wait_all_warp(L.wait, lane);
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair; t < L.total_tiles; t += num_pairs) {
        const TileCoord tc = decode_tile(L, t, BN, 2 * BM);
        const int kblocks = L.prob[tc.p].k / BK;
        const int m0 = tc.m0 + (int)cta * BM;
        const int n0 = tc.n0 + (int)cta * (BN / 2);
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
          uint8_t* sb = sa + Cfg::A_BYTES;
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * Cfg::STAGE_BYTES);
          const uint32_t bar = mapa_shared(smem_u32(&full_bar[stage]), 0);
          const int kidx = kb * BK;
          const int kc = L.prob[tc.p].k_chunk;
          tma_load_4d_pair(sa, &L.tmap_a[tc.p], bar, kidx % kc, m0, kidx / kc, tc.b);
          tma_load_2d_pair(sb, &L.tmap_b[tc.p], bar, kb * BK, n0);
          if (++stage == Cfg::STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (leader CTA only)
    if (leader && elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = pair; t < L.total_tiles; t += num_pairs) {
        const TileCoord tc = decode_tile(L, t, BN, 2 * BM);
        const int kblocks = L.prob[tc.p].k / BK;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * Cfg::ACC_STRIDE;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t sb = sa + Cfg::A_BYTES;
          const uint64_t da = make_smem_desc(sa, 0, 1024);
          const uint64_t db = make_smem_desc(sb, 0, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            mma_ss_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          tc_commit_pair(&empty_bar[stage], 3);
          if (++stage == Cfg::STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        tc_commit_pair(&tmem_full[acc], 3);
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue warps (both CTAs)
    const int q = warp & 3;
    uint8_t* epi_stage = smem + Cfg::STAGES * Cfg::STAGE_BYTES + (warp - 2) * Cfg::EPI_WARP_BYTES;   // 1024-aligned
    int acc = 0;
    uint32_t acc_phase = 0;
    const uint32_t empty0 = mapa_shared(smem_u32(&tmem_empty[0]), 0);
    const uint32_t empty1 = mapa_shared(smem_u32(&tmem_empty[1]), 0);
    for (int t = pair; t < L.total_tiles; t += num_pairs) {
      const TileCoord tc = decode_tile(L, t, BN, 2 * BM);
      const GemmProb& P = L.prob[tc.p];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * Cfg::ACC_STRIDE;
      epilogue_tile<BN>(P, tc.b, tc.m0 + (int)cta * BM + q * 32 + lane, tc.n0, taddr, (warp - 2) >> 2, epi_stage, [&]() {
        mbar_wait(&tmem_full[acc], acc_phase);
        tc_fence_after();
      }, [&]() {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(acc == 0 ? empty0 : empty1);
      }, L.sc_tma ? L.tmap_sc : nullptr);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (L.sc_tma && lane == 0) {
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
      asm volatile("fence.proxy.async;" ::: "memory");
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (threadIdx.x == 0) peer_notify_last_cta(L.notify, gridDim.x);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair<Cfg::TMEM_COLS>(tmem_base);
  }
}

}  // namespace qi

using namespace qi;

static int g_epi_dbg = 0;
extern "C" int qi_gemm_set_debug(int mask) {   // profiling aid (see the header)
  g_epi_dbg = mask;
  return QI_OK;
}

static int fill_problem(GemmLaunch& L, int i, const qi_gemm_problem_t& in, int BN, int tile_begin,
                        int tile_m = BM, int b_box_rows = 0) {
  if (!in.a || !in.w || (!in.c && !(in.epilogue == QI_EPI_QKV_NORM_ROPE && in.scatter))) return QI_ERR_ARG;
  if (in.m <= 0 || in.batch <= 0 || in.n <= 0 || in.k <= 0) return QI_ERR_ARG;
  if (in.k % BK != 0 || in.n % 8 != 0) return QI_ERR_UNSUPPORTED;
  if (in.lda % 8 != 0 || in.ldc % 8 != 0 || in.a_batch_stride % 8 != 0 || in.c_batch_stride % 8 != 0)
    return QI_ERR_ARG;
  if (in.epilogue < 0 || in.epilogue > 3) return QI_ERR_ARG;
  if (in.c_fp32 && in.epilogue != QI_EPI_BIAS && in.epilogue != QI_EPI_GATED_RESIDUAL) return QI_ERR_UNSUPPORTED;
  if (in.c_fp32 && in.epilogue == QI_EPI_GATED_RESIDUAL && (reinterpret_cast<uintptr_t>(in.residual) & 15))
    return QI_ERR_ARG;
  if (in.epilogue == QI_EPI_GATED_RESIDUAL &&
      (!in.residual || !in.gate || in.ld_res % 8 != 0 || in.res_batch_stride % 8 != 0 ||
       in.gate_batch_stride % 8 != 0))
    return QI_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(in.a) | reinterpret_cast<uintptr_t>(in.w) |
       reinterpret_cast<uintptr_t>(in.c)) & 15)
    return QI_ERR_ARG;
  GemmProb& P = L.prob[i];
  P.c = in.c;
  P.bias = (const __nv_bfloat16*)in.bias;
  P.residual = in.epilogue == QI_EPI_GATED_RESIDUAL ? (const __nv_bfloat16*)in.residual : nullptr;
  P.gate = in.epilogue == QI_EPI_GATED_RESIDUAL ? (const __nv_bfloat16*)in.gate : nullptr;
  P.ldc = in.ldc; P.c_bs = in.c_batch_stride;
  P.ld_res = in.ld_res; P.res_bs = in.res_batch_stride; P.gate_bs = in.gate_batch_stride;
  P.m = in.m; P.batch = in.batch; P.n = in.n; P.k = in.k;
  P.epilogue = in.epilogue;
  P.c_fp32 = in.c_fp32 ? 1 : 0;
  P.dbg = g_epi_dbg;
  P.gate_fp32 = (in.gate_fp32 && in.epilogue == QI_EPI_GATED_RESIDUAL) ? 1 : 0;
  if (P.gate_fp32 && ((reinterpret_cast<uintptr_t>(in.gate) & 15) || in.gate_batch_stride % 4 != 0)) return QI_ERR_ARG;
  P.freqs = nullptr; P.wq = P.wk = nullptr; P.sc_n = 0; P.sc_me = 0; P.sc_tpg = 0;
  if (in.epilogue == QI_EPI_QKV_NORM_ROPE) {
    if (!in.qk_norm_wq || !in.qk_norm_wk || !in.rope_cos_sin || in.rope_pos0 < 0) return QI_ERR_ARG;
    if (in.head_dim != 64 && in.head_dim != 128) return QI_ERR_UNSUPPORTED;
    if (in.heads <= 0 || (long long)in.heads * in.head_dim * 3 != in.n || BN % in.head_dim != 0) return QI_ERR_ARG;
    if ((reinterpret_cast<uintptr_t>(in.qk_norm_wq) | reinterpret_cast<uintptr_t>(in.qk_norm_wk) |
         reinterpret_cast<uintptr_t>(in.rope_cos_sin)) & 15)
      return QI_ERR_ARG;
    P.freqs = (const float2*)in.rope_cos_sin;
    P.wq = (const __nv_bfloat16*)in.qk_norm_wq;
    P.wk = (const __nv_bfloat16*)in.qk_norm_wk;
    P.pos0 = in.rope_pos0; P.heads = in.heads; P.head_dim = in.head_dim; P.qk_eps = in.qk_norm_eps;
    if (in.scatter) {
      const qi_peer_scatter_t& d = *in.scatter;
      if (d.n_peers < 1 || d.n_peers > 8 || d.my_index < 0 || d.my_index >= d.n_peers ||
          in.heads % d.n_peers != 0 || d.row_prefix[0] != 0)
        return QI_ERR_ARG;
      if (d.row_prefix[d.my_index] + in.rope_pos0 + in.m > d.row_prefix[d.my_index + 1]) return QI_ERR_ARG;
      for (int p = 0; p < d.n_peers; ++p) {
        if (!d.base[p] || (reinterpret_cast<uintptr_t>(d.base[p]) & 15)) return QI_ERR_ARG;
        P.sc_base[p] = (__nv_bfloat16*)d.base[p];
      }
      P.sc_n = d.n_peers;
      P.sc_hs = in.heads / d.n_peers;
      P.sc_s_total = d.row_prefix[d.n_peers];
      P.sc_row_offset = d.row_prefix[d.my_index];
      P.sc_me = d.my_index;
      static const bool remap = [] {
        const char* e = getenv("QI_SCATTER_REMAP");   // A/B aid: "0" keeps the plain column-tile order
        return !(e && e[0] == '0');
      }();
      const int owner_cols = P.sc_hs * in.head_dim;   // columns of one (q|k|v, owner) group
      if (remap && d.n_peers > 1 && owner_cols % BN == 0) P.sc_tpg = owner_cols / BN;
    }
  }
  P.tiles_m = (in.m + tile_m - 1) / tile_m;
  P.tiles_m_total = P.tiles_m * in.batch;
  P.tiles_n = (in.n + BN - 1) / BN;
  P.tile_begin = tile_begin;
  // A: [batch, k_chunks, m, k_chunk]  (innermost first for the tensor map)
  {
    const int kc = in.a_k_chunk > 0 ? in.a_k_chunk : in.k;
    if (kc % BK != 0 || in.k % kc != 0) return QI_ERR_ARG;
    if (in.a_k_chunk > 0 && in.a_k_chunk_stride % 8 != 0) return QI_ERR_ARG;
    P.k_chunk = kc;
    const int nchunks = in.k / kc;
    const uint64_t row_bytes = (uint64_t)in.lda * 2;
    uint64_t dims[4] = {(uint64_t)kc, (uint64_t)in.m, (uint64_t)nchunks, (uint64_t)in.batch};
    uint64_t strides[3] = {row_bytes,
                           nchunks > 1 ? (uint64_t)in.a_k_chunk_stride * 2 : row_bytes * (uint64_t)in.m,
                           in.batch > 1 ? (uint64_t)in.a_batch_stride * 2 : row_bytes * (uint64_t)in.m};
    uint32_t box[4] = {BK, BM, 1, 1};
    int rc = qi_encode_tmap_bf16(&L.tmap_a[i], in.a, 4, dims, strides, box);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)in.k, (uint64_t)in.n};
    uint64_t strides[1] = {(uint64_t)in.k * 2};
    uint32_t box[2] = {BK, (uint32_t)(b_box_rows ? b_box_rows : BN)};
    int rc = qi_encode_tmap_bf16(&L.tmap_b[i], in.w, 2, dims, strides, box);
    if (rc) return rc;
  }
  return QI_OK;
}

// Row tiles per supertile: ~sqrt(workers), so that one wave of `workers` tiles is a square patch of the tile grid
// (9 for 74 CTA pairs, 12 for 148 CTAs).  QI_GEMM_GROUP_M overrides (0 = column-major order, the round-1 schedule).
static int g_group_m = -1;   // -1: env QI_GEMM_GROUP_M or the default below
extern "C" int qi_gemm_set_group_m(int group_m) {
  if (group_m < -1 || group_m > 4096) return QI_ERR_ARG;
  g_group_m = group_m;
  return QI_OK;
}
static int gemm_group_m(int workers, int tuned = -1) {
  static const int env = [] {
    const char* e = getenv("QI_GEMM_GROUP_M");
    return e ? atoi(e) : -1;
  }();
  if (g_group_m >= 0) return g_group_m;
  if (tuned >= 0) return tuned;   // this launch's measured schedule
  if (env >= 0) return env;
  int g = 1;
  while ((g + 1) * (g + 1) <= workers) ++g;
  return g + ((g * (g + 1) <= workers) ? 1 : 0);   // round sqrt to nearest
}

// Scatter destinations as TMA-store tensors (QI_SCATTER_TMA=0: 16-byte st.global per lane instead).  Every problem
// of the launch must scatter into the same buffers (the img and txt streams of one block do).
static int fill_scatter_maps(GemmLaunch& L, const qi_gemm_problem_t* problems, int num) {
  static const bool on = [] {
    const char* e = getenv("QI_SCATTER_TMA");
    return !(e && e[0] == '0');
  }();
  const qi_peer_scatter_t* d = problems[0].scatter;
  if (!on || !d || problems[0].epilogue != QI_EPI_QKV_NORM_ROPE || problems[0].head_dim != 128) return QI_OK;
  for (int i = 1; i < num; ++i) {
    const qi_peer_scatter_t* e = problems[i].scatter;
    if (!e || e->n_peers != d->n_peers || problems[i].batch != problems[0].batch ||
        problems[i].heads != problems[0].heads || problems[i].head_dim != 128)
      return QI_OK;
    for (int p = 0; p < d->n_peers; ++p)
      if (e->base[p] != d->base[p]) return QI_OK;
  }
  const int hs = problems[0].heads / d->n_peers;
  const uint64_t s_total = (uint64_t)d->row_prefix[d->n_peers];
  const uint64_t dims[3] = {128, s_total, (uint64_t)problems[0].batch * 3 * hs};
  const uint64_t str[2] = {128 * 2, s_total * 128 * 2};
  const uint32_t box[3] = {64, 32, 1};
  for (int p = 0; p < d->n_peers; ++p)
    if (int rc = qi_encode_tmap_bf16(&L.tmap_sc[p], d->base[p], 3, dims, str, box)) return rc;
  L.sc_tma = 1;
  return QI_OK;
}

template <int BN>
static int launch_gemm(const qi_gemm_problem_t* problems, int num, cudaStream_t stream, int tuned_gm = -1) {
  GemmLaunch L;
  memset(&L, 0, sizeof(L));
  L.num_problems = num;
  int tiles = 0;
  for (int i = 0; i < num; ++i) {
    int rc = fill_problem(L, i, problems[i], BN, tiles);
    if (rc) return rc;
    tiles += L.prob[i].tiles_m_total * L.prob[i].tiles_n;
  }
  L.total_tiles = tiles;
  if (int rc = fill_signal(L.notify, problems[0].scatter ? problems[0].scatter->notify : nullptr)) return rc;
  if (int rc = fill_signal(L.wait, problems[0].wait)) return rc;
  if (int rc = fill_scatter_maps(L, problems, num)) return rc;
  if (int rc = qi_set_max_smem(reinterpret_cast<const void*>(gemm_bf16_kernel<BN>), GemmCfg<BN>::SMEM_BYTES)) return rc;
  const int sms = qi_num_sms();
  if (sms <= 0) return QI_ERR_CUDA;
  L.group_m = gemm_group_m(sms, tuned_gm);
  const int grid = tiles < sms ? tiles : sms;
  if (qi_launch(gemm_bf16_kernel<BN>, dim3(grid), dim3(GEMM_THREADS), GemmCfg<BN>::SMEM_BYTES, stream, L) !=
      cudaSuccess)
    return QI_ERR_CUDA;
  return QI_OK;
}

template <int BN, int STG = 0>
static int launch_gemm_pair(const qi_gemm_problem_t* problems, int num, cudaStream_t stream, int tuned_gm = -1) {
  GemmLaunch L;
  memset(&L, 0, sizeof(L));
  L.num_problems = num;
  int tiles = 0;
  for (int i = 0; i < num; ++i) {
    int rc = fill_problem(L, i, problems[i], BN, tiles, 2 * BM, BN / 2);
    if (rc) return rc;
    tiles += L.prob[i].tiles_m_total * L.prob[i].tiles_n;
  }
  L.total_tiles = tiles;
  if (int rc = fill_signal(L.notify, problems[0].scatter ? problems[0].scatter->notify : nullptr)) return rc;
  if (int rc = fill_signal(L.wait, problems[0].wait)) return rc;
  if (int rc = fill_scatter_maps(L, problems, num)) return rc;
  if (int rc = qi_set_max_smem(reinterpret_cast<const void*>(gemm_pair_bf16_kernel<BN, STG>), GemmPairCfg<BN, STG>::SMEM_BYTES))
    return rc;
  const int pairs_max = qi_num_sms() / 2;
  if (pairs_max <= 0) return QI_ERR_CUDA;
  L.group_m = gemm_group_m(pairs_max, tuned_gm);
  const int pairs = tiles < pairs_max ? tiles : pairs_max;
  if (qi_launch(gemm_pair_bf16_kernel<BN, STG>, dim3(2 * pairs), dim3(GEMM_THREADS),
                GemmPairCfg<BN, STG>::SMEM_BYTES, stream, L) != cudaSuccess)
    return QI_ERR_CUDA;
  return QI_OK;
}

static int g_force_bn = 0;
extern "C" int qi_gemm_set_tile_n(int bn) {
  if (bn != 0 && bn != 128 && bn != 160 && bn != 192 && bn != 224 && bn != 256) return QI_ERR_ARG;
  g_force_bn = bn;
  return QI_OK;
}

// Measured schedules: (problem dims) -> (pair, bn), filled by the host layer's autotuner
// (ops.tune_gemm: times every candidate once per new shape before the step is captured into a graph).
struct TunedSchedule {
  int num, m[2], batch[2], n[2], k[2], epi[2];   // epi = epilogue | c_fp32 << 8
  int pair, bn, group_m;   // group_m < 0: default rasterisation
};
static TunedSchedule g_tuned[512];
static int g_num_tuned = 0;
static bool tuned_matches(const TunedSchedule& t, const qi_gemm_problem_t* p, int num) {
  if (t.num != num) return false;
  for (int i = 0; i < num; ++i)
    if (t.m[i] != p[i].m || t.batch[i] != p[i].batch || t.n[i] != p[i].n || t.k[i] != p[i].k ||
        t.epi[i] != (p[i].epilogue | (p[i].c_fp32 ? 256 : 0)))
      return false;
  return true;
}
extern "C" int qi_gemm_set_schedule_ex(const qi_gemm_problem_t* problems, int num_problems, int pair, int bn,
                                       int group_m);
extern "C" int qi_gemm_set_schedule(const qi_gemm_problem_t* problems, int num_problems, int pair, int bn) {
  return qi_gemm_set_schedule_ex(problems, num_problems, pair, bn, -1);
}
extern "C" int qi_gemm_set_schedule_ex(const qi_gemm_problem_t* problems, int num_problems, int pair, int bn,
                                       int group_m) {
  if (group_m < -1 || group_m > 4096) return QI_ERR_ARG;
  if (!problems || num_problems < 1 || num_problems > 2 || pair < 0 || pair > 1) return QI_ERR_ARG;
  if (bn != 128 && bn != 160 && bn != 192 && bn != 224 && bn != 256) return QI_ERR_ARG;
  int slot = -1;
  for (int i = 0; i < g_num_tuned; ++i)
    if (tuned_matches(g_tuned[i], problems, num_problems)) slot = i;
  if (slot < 0) {
    if (g_num_tuned == 512) return QI_ERR_UNSUPPORTED;
    slot = g_num_tuned++;
  }
  TunedSchedule& t = g_tuned[slot];
  memset(&t, 0, sizeof(t));
  t.num = num_problems;
  for (int i = 0; i < num_problems; ++i) {
    t.m[i] = problems[i].m; t.batch[i] = problems[i].batch; t.n[i] = problems[i].n; t.k[i] = problems[i].k;
    t.epi[i] = problems[i].epilogue | (problems[i].c_fp32 ? 256 : 0);
  }
  t.pair = pair;
  t.bn = bn;
  t.group_m = group_m;
  return QI_OK;
}
extern "C" int qi_gemm_clear_schedules(void) {
  g_num_tuned = 0;
  return QI_OK;
}

static int g_pair_mode = -1;   // -1 auto, 0 never, 1 always
extern "C" int qi_gemm_set_cta_pair(int mode) {
  if (mode < -1 || mode > 1) return QI_ERR_ARG;
  g_pair_mode = mode;
  return QI_OK;
}

extern "C" int qi_gemm_bf16(const qi_gemm_problem_t* problems, int num_problems,
                            cudaStream_t stream) {
  if (!problems || num_problems < 1 || num_problems > 2) return QI_ERR_ARG;
  const int sms = qi_num_sms();
  if (sms <= 0) return QI_ERR_CUDA;
  // Schedule choice.  A persistent grid runs ceil(tiles / workers) waves; a tile costs ~BN (+ a fixed
  // per-tile overhead).  Candidates: single-CTA tiles 128 x {256,192,128} on 148 SMs and CTA-pair
  // tiles 256 x {256,192,128} on 74 pairs; the pair kernel moves a third less operand data per FLOP
  // (weight ~0.93 per unit of work) but quantises M in steps of 256.
  static const int env_pair = [] {
    const char* e = getenv("QI_GEMM_PAIR");   // debugging aid: "0" / "1" pins the kernel family
    return e ? atoi(e) : -1;
  }();
  const int pair_mode = g_pair_mode >= 0 ? g_pair_mode : (env_pair == 0 || env_pair == 1 ? env_pair : -1);
  int bn_multiple = 32;   // the QKV epilogue needs whole head vectors inside a tile
  for (int i = 0; i < num_problems; ++i)
    if (problems[i].epilogue == QI_EPI_QKV_NORM_ROPE && problems[i].head_dim > bn_multiple)
      bn_multiple = problems[i].head_dim;
  int bn = g_force_bn, pair = pair_mode, tuned_gm = -1;
  if (bn != 0 && bn % bn_multiple != 0) return QI_ERR_UNSUPPORTED;
  if (bn == 0 && pair < 0) {
    for (int i = 0; i < g_num_tuned; ++i)
      if (tuned_matches(g_tuned[i], problems, num_problems) && g_tuned[i].bn % bn_multiple == 0) {
        bn = g_tuned[i].bn;
        pair = g_tuned[i].pair;
        tuned_gm = g_tuned[i].group_m;
        break;
      }
  }
  if (bn == 0 || pair < 0) {
    double best = -1;
    const int cand[5] = {256, 224, 192, 160, 128};
    int best_bn = 256, best_pair = 0;
    for (int pm = 0; pm < 2; ++pm) {
      if (pair_mode >= 0 && pm != pair_mode) continue;
      for (int c = 0; c < 5; ++c) {
        if (g_force_bn && cand[c] != g_force_bn) continue;
        if (cand[c] % bn_multiple != 0) continue;
        const int tm = pm ? 2 * BM : BM;
        long long tiles = 0;
        for (int i = 0; i < num_problems; ++i) {
          const qi_gemm_problem_t& p = problems[i];
          tiles += (long long)((p.m + tm - 1) / tm) * p.batch * ((p.n + cand[c] - 1) / cand[c]);
        }
        const int workers = pm ? sms / 2 : sms;
        const long long waves = (tiles + workers - 1) / workers;
        const double cost = (double)waves * (cand[c] + 24) * (pm ? 0.93 : 1.0);
        if (best < 0 || cost < best) {
          best = cost;
          best_bn = cand[c];
          best_pair = pm;
        }
      }
    }
    bn = best_bn;
    pair = best_pair;
  }
  if (pair) {
    static const int env_stg = [] {
      const char* e = getenv("QI_GEMM_PAIR_STAGES");   // tuning aid (256-wide pair tile only)
      return e ? atoi(e) : 0;
    }();
    if (bn == 256 && env_stg == 3) return launch_gemm_pair<256, 3>(problems, num_problems, stream, tuned_gm);
    if (bn == 256 && env_stg == 4) return launch_gemm_pair<256, 4>(problems, num_problems, stream, tuned_gm);
    if (bn == 256 && env_stg == 5) return launch_gemm_pair<256, 5>(problems, num_problems, stream, tuned_gm);
    if (bn == 256) return launch_gemm_pair<256>(problems, num_problems, stream, tuned_gm);
    if (bn == 224) return launch_gemm_pair<224>(problems, num_problems, stream, tuned_gm);
    if (bn == 192) return launch_gemm_pair<192>(problems, num_problems, stream, tuned_gm);
    if (bn == 160) return launch_gemm_pair<160>(problems, num_problems, stream, tuned_gm);
    return launch_gemm_pair<128>(problems, num_problems, stream, tuned_gm);
  }
  if (bn == 256) return launch_gemm<256>(problems, num_problems, stream, tuned_gm);
  if (bn == 224) return launch_gemm<224>(problems, num_problems, stream, tuned_gm);
  if (bn == 192) return launch_gemm<192>(problems, num_problems, stream, tuned_gm);
  if (bn == 160) return launch_gemm<160>(problems, num_problems, stream, tuned_gm);
  return launch_gemm<128>(problems, num_problems, stream, tuned_gm);
}
