// Memory-bound kernels of the Qwen-Image denoising path (sm_100a):
//   K1  qi_adaln_modulate      LayerNorm(no affine) fused with adaLN shift/scale   (A.3)
//       qi_rmsnorm_affine      RMSNorm with learned weight (txt_norm)              (A.1)
//   K2  qi_qknorm_rope         per-head QK RMSNorm + MSRoPE rotation, in place     (A.4)
//   K5  qi_cfg_euler_step      true-CFG combine + norm rescale + FlowMatch Euler   (A.6)
//       qi_small_linear        M<=32 rows x [N,K] bf16 weights (time embed, modulation, norm_out)
//       qi_timestep_sinusoid   256-wide sinusoidal projection of t                 (A.2)
// All use 128-bit global accesses, one warp per row (or per group of heads/tokens), and
// warp-shuffle reductions; no shared memory is needed because no datum is reused across warps.
#include "qi_common.cuh"
#include "../../include/qi_b200.h"

namespace qi {

__device__ __forceinline__ uint4 ldg128(const void* p) {
  return *reinterpret_cast<const uint4*>(p);
}
__device__ __forceinline__ uint4 ldg128_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg128(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }
__device__ __forceinline__ uint2 ldg64(const void* p) { return *reinterpret_cast<const uint2*>(p); }
__device__ __forceinline__ void stg64(void* p, uint2 v) { *reinterpret_cast<uint2*>(p) = v; }
__device__ __forceinline__ float4 ldg_f4_stream(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ void unpack8(uint4 u, float (&f)[8]) {
  float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), c = unpack_bf16(u.z), d = unpack_bf16(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]),
                    pack_bf16(f[6], f[7]));
}
template <int W>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------------------------------------
// K1: one warp per row.  NCHUNK = d / 256 (each lane owns NCHUNK x 8 contiguous bf16).
// MODE 0: y = LN(x) * (1 + scale[b]) + shift[b]     MODE 1: y = bf16(x * rsqrt(ms+eps)) * w
// ---------------------------------------------------------------------------------------------
// Block-per-row variant: 128 threads own one row, CPT 16-byte chunks per thread (d = CPT * 1024).
// ~40 registers per thread, so an SM keeps 16 rows (~100 KB of loads) in flight; the two
// reductions (mean, then centred variance) go warp shuffle -> 4 partials in shared memory.
struct NormSeg {   // one set of rows sharing strides and modulation vectors
  const __nv_bfloat16* x;    // fp32 rows when the kernel is instantiated with XF32 (the fp32 residual streams)
  __nv_bfloat16* y;
  const __nv_bfloat16* shift;
  const __nv_bfloat16* scale;
  long long mod_batch_stride;
  int rows, ld_x, ld_y, rows_per_batch;
  int mod_fp32;              // shift / scale are fp32 vectors (the fp32 modulation table), else bf16
};

// Up to two segments per launch (the img and txt streams of a block): block b < seg0.rows works on
// segment 0, the others on segment 1.
// XF32: x is fp32 (the residual streams h / e are kept in fp32).  A thread then owns, per 1024-element
// chunk, elements [4 tid, +4) and [512 + 4 tid, +4): two fully coalesced 16-byte loads in, two 8-byte stores out.
// 8 fp32 modulation values for the elements a thread owns in chunk c (same ownership as the x loads)
template <bool XF32>
__device__ __forceinline__ void load_mod_f32(const float* p, int c, int tid, float (&o)[8]) {
  float4 a, b;
  if constexpr (XF32) {
    a = *reinterpret_cast<const float4*>(p + c * 1024 + tid * 4);
    b = *reinterpret_cast<const float4*>(p + c * 1024 + 512 + tid * 4);
  } else {
    const float4* q = reinterpret_cast<const float4*>(p + (c * 128 + tid) * 8);
    a = q[0];
    b = q[1];
  }
  o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}

template <int CPT, int MODE, bool XF32>
__global__ void __launch_bounds__(128) norm_rows_kernel(const __grid_constant__ NormSeg s0,
                                                        const __grid_constant__ NormSeg s1,
                                                        float eps) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[2][4];
  const bool second = blockIdx.x >= s0.rows;
  const int row = second ? blockIdx.x - s0.rows : blockIdx.x;
  const __nv_bfloat16* x = second ? s1.x : s0.x;
  __nv_bfloat16* y = second ? s1.y : s0.y;
  const __nv_bfloat16* shift = second ? s1.shift : s0.shift;
  const __nv_bfloat16* scale = second ? s1.scale : s0.scale;
  const long long mod_batch_stride = second ? s1.mod_batch_stride : s0.mod_batch_stride;
  const int ld_x = second ? s1.ld_x : s0.ld_x, ld_y = second ? s1.ld_y : s0.ld_y;
  const int rows_per_batch = second ? s1.rows_per_batch : s0.rows_per_batch;
  const int tid = threadIdx.x;
  const int lane = tid & 31, wid = tid >> 5;
  constexpr int D = CPT * 1024;
  float v[CPT][8];
  if constexpr (XF32) {
    const float* xr = reinterpret_cast<const float*>(x) + (size_t)row * ld_x;
    float4 raw[CPT][2];
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      raw[c][0] = ldg_f4_stream(xr + c * 1024 + tid * 4);
      raw[c][1] = ldg_f4_stream(xr + c * 1024 + 512 + tid * 4);
    }
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      v[c][0] = raw[c][0].x; v[c][1] = raw[c][0].y; v[c][2] = raw[c][0].z; v[c][3] = raw[c][0].w;
      v[c][4] = raw[c][1].x; v[c][5] = raw[c][1].y; v[c][6] = raw[c][1].z; v[c][7] = raw[c][1].w;
    }
  } else {
    const __nv_bfloat16* xr = x + (size_t)row * ld_x;
    uint4 raw[CPT];
#pragma unroll
    for (int c = 0; c < CPT; ++c) raw[c] = ldg128_stream(xr + (c * 128 + tid) * 8);
#pragma unroll
    for (int c = 0; c < CPT; ++c) unpack8(raw[c], v[c]);
  }
  // modulation vectors do not depend on the row statistics: issue their loads now
  const int b = row / rows_per_batch;
  const __nv_bfloat16* sh = (MODE == 0) ? shift + (size_t)b * mod_batch_stride : nullptr;
  const __nv_bfloat16* sc = scale + ((MODE == 0) ? (size_t)b * mod_batch_stride : 0);
  // fp32 modulation vectors (12 KB per batch item, L1/L2 resident) are read after the reductions instead, which
  // keeps the register count (and with it the rows in flight per SM) where the bf16 variant has it
  const bool mf32 = (MODE == 0) && s0.mod_fp32;
  const float* sh32 = reinterpret_cast<const float*>(shift) + (size_t)b * mod_batch_stride;
  const float* sc32 = reinterpret_cast<const float*>(scale) + (size_t)b * mod_batch_stride;
  uint4 wraw[CPT], hraw[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    if (mf32) break;
    if constexpr (XF32) {
      const uint2 lo = ldg64(sc + c * 1024 + tid * 4), hi = ldg64(sc + c * 1024 + 512 + tid * 4);
      wraw[c] = make_uint4(lo.x, lo.y, hi.x, hi.y);
      if (MODE == 0) {
        const uint2 l2 = ldg64(sh + c * 1024 + tid * 4), h2 = ldg64(sh + c * 1024 + 512 + tid * 4);
        hraw[c] = make_uint4(l2.x, l2.y, h2.x, h2.y);
      }
    } else {
      wraw[c] = ldg128(sc + (c * 128 + tid) * 8);
      if (MODE == 0) hraw[c] = ldg128(sh + (c * 128 + tid) * 8);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < CPT; ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) s += (MODE == 0) ? v[c][j] : v[c][j] * v[c][j];
  s = group_sum<32>(s);
  if (lane == 0) red[0][wid] = s;
  __syncthreads();
  s = red[0][0] + red[0][1] + red[0][2] + red[0][3];
  float mean = 0.f, rstd;
  if (MODE == 0) {
    mean = s * (1.f / D);
    float q = 0.f;
#pragma unroll
    for (int c = 0; c < CPT; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float t = v[c][j] - mean;
        q = fmaf(t, t, q);
      }
    q = group_sum<32>(q);
    if (lane == 0) red[1][wid] = q;
    __syncthreads();
    q = red[1][0] + red[1][1] + red[1][2] + red[1][3];
    rstd = rsqrtf(q * (1.f / D) + eps);
  } else {
    rstd = rsqrtf(s * (1.f / D) + eps);
  }
  __nv_bfloat16* yr = y + (size_t)row * ld_y;
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    float w[8], o[8];
    if (mf32) load_mod_f32<XF32>(sc32, c, tid, w);
    else unpack8(wraw[c], w);
    if (MODE == 0) {
      float h[8];
      if (mf32) load_mod_f32<XF32>(sh32, c, tid, h);
      else unpack8(hraw[c], h);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[c][j] - mean) * rstd * (1.f + w[j]) + h[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = bf16_round(v[c][j] * rstd) * w[j];
    }
    const uint4 pk = pack8(o);
    if constexpr (XF32) {
      stg64(yr + c * 1024 + tid * 4, make_uint2(pk.x, pk.y));
      stg64(yr + c * 1024 + 512 + tid * 4, make_uint2(pk.z, pk.w));
    } else {
      stg128(yr + (c * 128 + tid) * 8, pk);
    }
  }
}


// Generic-width variant (any d % 8 == 0): the row is re-read from L1/L2 for the second and third
// pass instead of being held in registers.  Only used for widths without a register instantiation
// (e.g. the tiny test config's joint_attention_dim = 128).
template <bool XF32>
__device__ __forceinline__ void load_row8(const __nv_bfloat16* xr, int off, float (&v)[8]) {
  if constexpr (XF32) {
    const float4* p = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(xr) + off);
    const float4 a = p[0], b = p[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
    unpack8(ldg128(xr + off), v);
  }
}

template <int MODE, bool XF32>
__global__ void __launch_bounds__(128) norm_rows_generic_kernel(const __grid_constant__ NormSeg s0,
                                                                const __grid_constant__ NormSeg s1,
                                                                int d, float eps) {
  pdl_launch_dependents();
  pdl_wait();
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= s0.rows + s1.rows) return;
  const bool second = warp >= s0.rows;
  if (second) warp -= s0.rows;
  const __nv_bfloat16* x = second ? s1.x : s0.x;
  __nv_bfloat16* y = second ? s1.y : s0.y;
  const __nv_bfloat16* shift = second ? s1.shift : s0.shift;
  const __nv_bfloat16* scale = second ? s1.scale : s0.scale;
  const long long mod_batch_stride = second ? s1.mod_batch_stride : s0.mod_batch_stride;
  const int ld_x = second ? s1.ld_x : s0.ld_x, ld_y = second ? s1.ld_y : s0.ld_y;
  const int rows_per_batch = second ? s1.rows_per_batch : s0.rows_per_batch;
  // (for XF32 the row pointer is kept in bf16 units of the fp32 row: load_row8 re-derives the float address)
  const __nv_bfloat16* xr = XF32 ? reinterpret_cast<const __nv_bfloat16*>(reinterpret_cast<const float*>(x) + (size_t)warp * ld_x)
                                 : x + (size_t)warp * ld_x;
  float s = 0.f;
  for (int off = lane * 8; off < d; off += 256) {
    float v[8];
    load_row8<XF32>(xr, off, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) s += (MODE == 0) ? v[j] : v[j] * v[j];
  }
  s = group_sum<32>(s);
  float mean = 0.f, rstd;
  if (MODE == 0) {
    mean = s / d;
    float q = 0.f;
    for (int off = lane * 8; off < d; off += 256) {
      float v[8];
      load_row8<XF32>(xr, off, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) q += (v[j] - mean) * (v[j] - mean);
    }
    q = group_sum<32>(q);
    rstd = rsqrtf(q / d + eps);
  } else {
    rstd = rsqrtf(s / d + eps);
  }
  const int b = warp / rows_per_batch;
  const __nv_bfloat16* sh = (MODE == 0) ? shift + (size_t)b * mod_batch_stride : nullptr;
  const __nv_bfloat16* sc = scale + ((MODE == 0) ? (size_t)b * mod_batch_stride : 0);
  __nv_bfloat16* yr = y + (size_t)warp * ld_y;
  for (int off = lane * 8; off < d; off += 256) {
    float v[8], w[8], o[8];
    load_row8<XF32>(xr, off, v);
    const bool mf32 = (MODE == 0) && s0.mod_fp32;
    if (mf32) load_row8<true>(reinterpret_cast<const __nv_bfloat16*>(reinterpret_cast<const float*>(scale) + (size_t)b * mod_batch_stride), off, w);
    else unpack8(ldg128(sc + off), w);
    if (MODE == 0) {
      float h[8];
      if (mf32) load_row8<true>(reinterpret_cast<const __nv_bfloat16*>(reinterpret_cast<const float*>(shift) + (size_t)b * mod_batch_stride), off, h);
      else unpack8(ldg128(sh + off), h);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[j] - mean) * rstd * (1.f + w[j]) + h[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = bf16_round(v[j] * rstd) * w[j];
    }
    stg128(yr + off, pack8(o));
  }
}

// ---------------------------------------------------------------------------------------------
// K2: QK RMSNorm + RoPE in place on the joint qkv buffer [rows, ld] (q at col 0, k at col d).
// A group of DH/8 lanes owns one head vector; a warp owns 32/(DH/8) consecutive head vectors.
// Rounding order follows upstream: bf16(x*rstd) -> bf16(*w) -> fp32 complex rotate -> bf16.
// ---------------------------------------------------------------------------------------------
template <int DH>
__global__ void __launch_bounds__(256) qknorm_rope_kernel(
    __nv_bfloat16* __restrict__ qkv, int rows, int ld, int heads, int seq, int s_txt,
    const __nv_bfloat16* __restrict__ wq_img, const __nv_bfloat16* __restrict__ wk_img,
    const __nv_bfloat16* __restrict__ wq_txt, const __nv_bfloat16* __restrict__ wk_txt,
    const float2* __restrict__ freqs /* [seq, DH/2] (cos, sin), joint order txt first */,
    float eps) {
  pdl_launch_dependents();
  pdl_wait();
  // One warp per token row.  A group of LPH = DH/8 lanes owns one head vector (8 bf16 per lane);
  // the warp walks the row's 2*heads head vectors (q heads then k heads, contiguous in memory)
  // VPW at a time.  The row's rotation table and the norm weights are loaded once per row.
  constexpr int LPH = DH / 8;
  constexpr int VPW = 32 / LPH;
  const int lane = threadIdx.x & 31;
  const int sub = lane % LPH;
  const int grp = lane / LPH;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  const int pos = row % seq;
  const bool is_txt = pos < s_txt;
  const float4* fp = reinterpret_cast<const float4*>(freqs + (size_t)pos * (DH / 2) + sub * 4);
  const float4 f01 = fp[0], f23 = fp[1];
  const float cs[4] = {f01.x, f01.z, f23.x, f23.z};
  const float sn[4] = {f01.y, f01.w, f23.y, f23.w};
  float wq[8], wk[8];
  unpack8(ldg128((is_txt ? wq_txt : wq_img) + sub * 8), wq);
  unpack8(ldg128((is_txt ? wk_txt : wk_img) + sub * 8), wk);
  __nv_bfloat16* base = qkv + (size_t)row * ld + sub * 8;
  const int nvec = 2 * heads;
  constexpr int UNROLL = 4;
  for (int v0 = grp; v0 < nvec; v0 += VPW * UNROLL) {
    uint4 raw[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int vi = v0 + u * VPW;
      if (vi < nvec) raw[u] = ldg128(base + (size_t)vi * DH);
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int vi = v0 + u * VPW;
      // all lanes of a group share vi, so the shuffles inside group_sum stay convergent per group;
      // groups past the end still execute them on stale registers and simply skip the store
      float v[8];
      unpack8(raw[u], v);
      float ss = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) ss = fmaf(v[j], v[j], ss);
      ss = group_sum<LPH>(ss);
      const float rstd = rsqrtf(ss * (1.f / DH) + eps);
      const bool is_k = vi >= heads;
      float o[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float a = bf16_round(bf16_round(v[2 * j] * rstd) * (is_k ? wk[2 * j] : wq[2 * j]));
        const float bb = bf16_round(bf16_round(v[2 * j + 1] * rstd) * (is_k ? wk[2 * j + 1] : wq[2 * j + 1]));
        o[2 * j] = a * cs[j] - bb * sn[j];
        o[2 * j + 1] = a * sn[j] + bb * cs[j];
      }
      if (vi < nvec) stg128(base + (size_t)vi * DH, pack8(o));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// K2 + Ulysses scatter: same arithmetic, but the result is not written back in place: head vector
// (row, j in {q,k,v}, head h) goes straight to the rank that owns head h, i.e. into that rank's
// HEAD-MAJOR [batch, 3, heads/P, s_total, DH] attention input buffer (peer-mapped over NVLink), at
// sequence row row_offset + s.  v vectors are moved untouched.  This IS the head<->sequence all-to-all.
// Head-major: consecutive tokens of one head are adjacent at the destination (256-byte row stride instead
// of 3*heads/P*DH*2 bytes) — measured peer write rate 574 vs 324 GB/s for 128-byte segments (tools/p2p_bw.py).
// ---------------------------------------------------------------------------------------------
struct ScatterDst {
  __nv_bfloat16* base[8];
  int n;               // ranks
  int heads_per_rank;
  int s_total;         // rows per batch item in the destination buffers
  int row_offset;      // first destination row of this rank's tokens
};

template <int DH>
__global__ void __launch_bounds__(256) qknorm_rope_scatter_kernel(
    const __nv_bfloat16* __restrict__ qkv, int rows, int ld, int heads, int seq, int s_txt,
    const __nv_bfloat16* __restrict__ wq_img, const __nv_bfloat16* __restrict__ wk_img,
    const __nv_bfloat16* __restrict__ wq_txt, const __nv_bfloat16* __restrict__ wk_txt,
    const float2* __restrict__ freqs, float eps, const __grid_constant__ ScatterDst D) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int LPH = DH / 8;
  constexpr int VPW = 32 / LPH;
  const int lane = threadIdx.x & 31;
  const int sub = lane % LPH;
  const int grp = lane / LPH;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  const int b = row / seq;
  const int pos = row - b * seq;
  const bool is_txt = pos < s_txt;
  const float4* fp = reinterpret_cast<const float4*>(freqs + (size_t)pos * (DH / 2) + sub * 4);
  const float4 f01 = fp[0], f23 = fp[1];
  const float cs[4] = {f01.x, f01.z, f23.x, f23.z};
  const float sn[4] = {f01.y, f01.w, f23.y, f23.w};
  float wq[8], wk[8];
  unpack8(ldg128((is_txt ? wq_txt : wq_img) + sub * 8), wq);
  unpack8(ldg128((is_txt ? wk_txt : wk_img) + sub * 8), wk);
  const __nv_bfloat16* src = qkv + (size_t)row * ld + sub * 8;
  const int hs = D.heads_per_rank;
  // destination (in units of DH-vectors): ((b*3 + j)*hs + hl)*s_total + row_offset + pos
  const size_t dpos = (size_t)D.row_offset + pos;
  const int nvec = 3 * heads;
  constexpr int UNROLL = 4;
  for (int v0 = grp; v0 < nvec; v0 += VPW * UNROLL) {
    uint4 raw[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int vi = v0 + u * VPW;
      if (vi < nvec) raw[u] = ldg128_stream(src + (size_t)vi * DH);
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int vi = v0 + u * VPW;
      float v[8];
      unpack8(raw[u], v);
      float ss = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) ss = fmaf(v[j], v[j], ss);
      ss = group_sum<LPH>(ss);
      const float rstd = rsqrtf(ss * (1.f / DH) + eps);
      const int j3 = vi / heads;            // 0 q, 1 k, 2 v
      const int h = vi - j3 * heads;
      const bool is_k = j3 == 1;
      float o[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float a = bf16_round(bf16_round(v[2 * j] * rstd) * (is_k ? wk[2 * j] : wq[2 * j]));
        const float bb = bf16_round(bf16_round(v[2 * j + 1] * rstd) * (is_k ? wk[2 * j + 1] : wq[2 * j + 1]));
        o[2 * j] = a * cs[j] - bb * sn[j];
        o[2 * j + 1] = a * sn[j] + bb * cs[j];
      }
      if (vi < nvec) {
        const int p = h / hs;
        const int hl = h - p * hs;
        __nv_bfloat16* dst = D.base[p] + ((((size_t)b * 3 + j3) * hs + hl) * D.s_total + dpos) * DH + sub * 8;
        stg128(dst, j3 == 2 ? raw[u] : pack8(o));
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// K5: one group of 8 lanes per token (64 channels = 8 lanes x 8 bf16).  Rounding follows the
// eager bf16 pipeline op by op: comb, the two norms, their ratio, the product,
// dt*v (dt itself rounded to bf16 by type promotion), then the fp32 accumulate and final cast.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cfg_euler_kernel(
    const __nv_bfloat16* __restrict__ cond, const __nv_bfloat16* __restrict__ uncond,
    const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, long long tokens,
    float cfg_scale, const float* __restrict__ dt_ptr, float dt_scalar) {
  pdl_launch_dependents();
  pdl_wait();
  const long long tok = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int sub = threadIdx.x & 7;
  const bool active = tok < tokens;
  const size_t off = (size_t)(active ? tok : tokens - 1) * 64 + sub * 8;
  const float dt = bf16_round(dt_ptr ? *dt_ptr : dt_scalar);
  float c[8], xv[8], v[8];
  unpack8(ldg128_stream(cond + off), c);
  unpack8(ldg128_stream(x + off), xv);
  if (uncond != nullptr) {
    float u[8], comb[8];
    unpack8(ldg128_stream(uncond + off), u);
    float cn = 0.f, nn = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      comb[j] = bf16_round(u[j] + bf16_round(cfg_scale * bf16_round(c[j] - u[j])));
      cn += c[j] * c[j];
      nn += comb[j] * comb[j];
    }
    cn = group_sum<8>(cn);
    nn = group_sum<8>(nn);
    const float ratio = bf16_round(bf16_round(sqrtf(cn)) / bf16_round(sqrtf(nn)));
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = bf16_round(comb[j] * ratio);
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = c[j];
  }
  float o[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j] = xv[j] + bf16_round(dt * v[j]);
  if (active) stg128(out + off, pack8(o));
}

// K5, fp32 form: predictions, latents and all arithmetic in fp32 (the form the pipeline uses).  True CFG amplifies
// whatever error the two predictions carry by (2 s - 1) = 7 at s = 4 before the norm rescale, and a bf16 latent
// re-rounded every step random-walks away from the fp32 trajectory (sqrt(50) * 1.7e-3 = 1.2e-2 over 50 steps),
// so the loop state is an fp32 master copy; `out_bf16` is the copy the next forward's img_in GEMM reads.
// 16 lanes per token (64 channels = 16 lanes x 4 fp32).
__global__ void __launch_bounds__(256) cfg_euler_f32_kernel(
    const float* __restrict__ cond, const float* __restrict__ uncond, const float* x /* may alias out */,
    float* out, __nv_bfloat16* __restrict__ out_bf16, long long tokens, float cfg_scale,
    const float* __restrict__ dt_ptr, float dt_scalar) {
  pdl_launch_dependents();
  pdl_wait();
  const long long tok = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
  const int sub = threadIdx.x & 15;
  const bool active = tok < tokens;
  const size_t off = (size_t)(active ? tok : tokens - 1) * 64 + sub * 4;
  const float dt = dt_ptr ? *dt_ptr : dt_scalar;
  const float4 c = *reinterpret_cast<const float4*>(cond + off);
  const float4 xv = *reinterpret_cast<const float4*>(x + off);
  float4 v = c;
  if (uncond != nullptr) {
    const float4 u = *reinterpret_cast<const float4*>(uncond + off);
    const float4 comb = make_float4(fmaf(cfg_scale, c.x - u.x, u.x), fmaf(cfg_scale, c.y - u.y, u.y),
                                    fmaf(cfg_scale, c.z - u.z, u.z), fmaf(cfg_scale, c.w - u.w, u.w));
    float cn = c.x * c.x + c.y * c.y + c.z * c.z + c.w * c.w;
    float nn = comb.x * comb.x + comb.y * comb.y + comb.z * comb.z + comb.w * comb.w;
    cn = group_sum<16>(cn);
    nn = group_sum<16>(nn);
    const float ratio = sqrtf(cn) / sqrtf(nn);
    v = make_float4(comb.x * ratio, comb.y * ratio, comb.z * ratio, comb.w * ratio);
  }
  const float4 o = make_float4(fmaf(dt, v.x, xv.x), fmaf(dt, v.y, xv.y), fmaf(dt, v.z, xv.z), fmaf(dt, v.w, xv.w));
  if (active) {
    *reinterpret_cast<float4*>(out + off) = o;
    if (out_bf16 != nullptr) stg64(out_bf16 + off, make_uint2(pack_bf16(o.x, o.y), pack_bf16(o.z, o.w)));
  }
}

// ---------------------------------------------------------------------------------------------
// Small-M linear: y[m, n] = act_out( sum_k act_in(x[m,k]) * W[n,k] + bias[n] ), M <= 32.
// One warp per output column n: the weight row is streamed once (128-bit loads) and reused for
// every m from registers; x stays L1/L2 resident.  HBM-bound on the weight bytes.
// ---------------------------------------------------------------------------------------------
// F32: x and y are fp32 and nothing is rounded in between (the timestep embedder and the modulation Linear feed
// every block of every step: their bf16 rounding is a coherent error, the largest single term of the path's
// distance to the fp32 oracle — tools/rounding_ablation.py).  Weights stay bf16.
template <int MT, bool F32>
__global__ void __launch_bounds__(256) small_linear_kernel(
    const void* __restrict__ x_, const __nv_bfloat16* __restrict__ W,
    const __nv_bfloat16* __restrict__ bias, void* __restrict__ y_, int M, int N, int K,
    int ld_x, int ld_y, int silu_in, int silu_out) {
  const __nv_bfloat16* x = reinterpret_cast<const __nv_bfloat16*>(x_);
  __nv_bfloat16* y = reinterpret_cast<__nv_bfloat16*>(y_);
  pdl_launch_dependents();
  pdl_wait();
  const int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  const int m0 = blockIdx.y * MT;
  float acc[MT];
#pragma unroll
  for (int m = 0; m < MT; ++m) acc[m] = 0.f;
  const __nv_bfloat16* wr = W + (size_t)n * K;
  for (int k = lane * 8; k < K; k += 256) {
    float w[8];
    unpack8(ldg128_stream(wr + k), w);
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      if (m0 + m < M) {
        float xv[8];
        if constexpr (F32) {
          const float4* xp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(x_) + (size_t)(m0 + m) * ld_x + k);
          const float4 a4 = xp[0], b4 = xp[1];
          xv[0] = a4.x; xv[1] = a4.y; xv[2] = a4.z; xv[3] = a4.w; xv[4] = b4.x; xv[5] = b4.y; xv[6] = b4.z; xv[7] = b4.w;
        } else {
          unpack8(ldg128(x + (size_t)(m0 + m) * ld_x + k), xv);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float a = xv[j];
          if (silu_in) a = F32 ? a / (1.f + expf(-a)) : bf16_round(a / (1.f + __expf(-a)));
          acc[m] = fmaf(a, w[j], acc[m]);
        }
      }
    }
  }
  const float bv = bias ? __bfloat162float(bias[n]) : 0.f;
#pragma unroll
  for (int m = 0; m < MT; ++m) {
    float s = group_sum<32>(acc[m]);
    if (lane == 0 && m0 + m < M) {
      s += bv;
      if (silu_out) {
        if (!F32) s = bf16_round(s);
        s = F32 ? s / (1.f + expf(-s)) : s / (1.f + __expf(-s));
      }
      if constexpr (F32) reinterpret_cast<float*>(y_)[(size_t)(m0 + m) * ld_y + n] = s;
      else y[(size_t)(m0 + m) * ld_y + n] = __float2bfloat16_rn(s);
    }
  }
}

// t [M] fp32 (already rounded the way the caller wants) -> [M, 256] bf16 = [cos(1000 t f_k), sin]
template <bool F32>
__global__ void timestep_sinusoid_kernel(const float* __restrict__ t, void* __restrict__ out_, int M) {
  pdl_launch_dependents();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M * 128) return;
  const int m = i / 128, k = i % 128;
  const float f = expf(-9.210340371976184f * (float)k / 128.f);
  const float a = 1000.f * (t[m] * f);
  if constexpr (F32) {
    float* out = reinterpret_cast<float*>(out_);
    out[(size_t)m * 256 + k] = cosf(a);
    out[(size_t)m * 256 + 128 + k] = sinf(a);
  } else {
    __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(out_);
    out[(size_t)m * 256 + k] = __float2bfloat16_rn(cosf(a));
    out[(size_t)m * 256 + 128 + k] = __float2bfloat16_rn(sinf(a));
  }
}

}  // namespace qi

// =============================================================================================
// C ABI
// =============================================================================================
using namespace qi;

#define QI_CHECK_LAUNCH()                              \
  do {                                                 \
    cudaError_t e__ = cudaGetLastError();              \
    if (e__ != cudaSuccess) return QI_ERR_CUDA;        \
  } while (0)

template <int MODE, bool XF32 = false>
static int launch_norm_rows(const NormSeg& a, const NormSeg& b, int d, float eps, cudaStream_t st) {
  const int rows = a.rows + b.rows;
  if (rows <= 0) return QI_OK;
  if (d % 8 != 0 || a.ld_x % 8 != 0 || a.ld_y % 8 != 0 || (b.rows > 0 && (b.ld_x % 8 != 0 || b.ld_y % 8 != 0)))
    return QI_ERR_ARG;
  const int blocks_generic = (rows + 3) / 4;
#define QI_NORM_CASE(NC)                                            \
  case NC:                                                          \
    qi_launch(norm_rows_kernel<NC, MODE, XF32>, dim3(rows), dim3(128), 0, st, a, b, eps); \
    break;
  switch (d % 1024 == 0 ? d / 1024 : 0) {
    QI_NORM_CASE(1) QI_NORM_CASE(2) QI_NORM_CASE(3) QI_NORM_CASE(4)
    default:
      qi_launch(norm_rows_generic_kernel<MODE, XF32>, dim3(blocks_generic), dim3(128), 0, st, a, b, d, eps);
  }
#undef QI_NORM_CASE
  QI_CHECK_LAUNCH();
  return QI_OK;
}

static NormSeg make_seg(const void* x, void* y, int rows, int ld_x, int ld_y, const void* shift,
                        const void* scale, long long mod_stride, int rows_per_batch) {
  NormSeg s;
  s.x = (const __nv_bfloat16*)x;
  s.y = (__nv_bfloat16*)y;
  s.shift = (const __nv_bfloat16*)shift;
  s.scale = (const __nv_bfloat16*)scale;
  s.mod_batch_stride = mod_stride;
  s.rows = rows;
  s.ld_x = ld_x;
  s.ld_y = ld_y;
  s.rows_per_batch = rows_per_batch;
  s.mod_fp32 = 0;
  return s;
}

extern "C" int qi_adaln_modulate(const void* x, void* y, int rows, int d, int ld_x, int ld_y,
                                 const void* shift, const void* scale, long long mod_batch_stride,
                                 int rows_per_batch, float eps, cudaStream_t stream) {
  if (!x || !y || !shift || !scale || rows_per_batch <= 0) return QI_ERR_ARG;
  NormSeg none;
  memset(&none, 0, sizeof(none));
  return launch_norm_rows<0>(make_seg(x, y, rows, ld_x, ld_y, shift, scale, mod_batch_stride, rows_per_batch),
                             none, d, eps, stream);
}

extern "C" int qi_adaln_modulate_multi(const qi_norm_rows_t* segs, int num_segs, int d, float eps,
                                       cudaStream_t stream) {
  if (!segs || num_segs < 1 || num_segs > 2) return QI_ERR_ARG;
  NormSeg s[2];
  memset(s, 0, sizeof(s));
  s[1].mod_fp32 = segs[0].mod_fp32 ? 1 : 0;
  for (int i = 0; i < num_segs; ++i) {
    const qi_norm_rows_t& g = segs[i];
    if (!g.x || !g.y || !g.shift || !g.scale || g.rows_per_batch <= 0 || g.rows < 0) return QI_ERR_ARG;
    if ((g.x_fp32 != 0) != (segs[0].x_fp32 != 0) || (g.mod_fp32 != 0) != (segs[0].mod_fp32 != 0))
      return QI_ERR_ARG;   // one input type and one modulation type per launch
    if (g.mod_fp32 && ((reinterpret_cast<uintptr_t>(g.shift) | reinterpret_cast<uintptr_t>(g.scale)) & 15 ||
                       g.mod_batch_stride % 4 != 0))
      return QI_ERR_ARG;
    s[i] = make_seg(g.x, g.y, g.rows, g.ld_x, g.ld_y, g.shift, g.scale, g.mod_batch_stride, g.rows_per_batch);
    s[i].mod_fp32 = g.mod_fp32 ? 1 : 0;
  }
  if (segs[0].x_fp32) return launch_norm_rows<0, true>(s[0], s[1], d, eps, stream);
  return launch_norm_rows<0>(s[0], s[1], d, eps, stream);
}

extern "C" int qi_rmsnorm_affine(const void* x, void* y, int rows, int d, int ld_x, int ld_y,
                                 const void* weight, float eps, cudaStream_t stream) {
  if (!x || !y || !weight) return QI_ERR_ARG;
  NormSeg none;
  memset(&none, 0, sizeof(none));
  return launch_norm_rows<1>(make_seg(x, y, rows, ld_x, ld_y, nullptr, weight, 0, 1 << 30), none, d, eps,
                             stream);
}

extern "C" int qi_qknorm_rope(void* qkv, int rows, int ld, int heads, int head_dim, int seq,
                              int s_txt, const void* wq_img, const void* wk_img,
                              const void* wq_txt, const void* wk_txt, const void* freqs_cos_sin,
                              float eps, cudaStream_t stream) {
  if (!qkv || !wq_img || !wk_img || !wq_txt || !wk_txt || !freqs_cos_sin) return QI_ERR_ARG;
  if (rows <= 0) return QI_OK;
  if (ld % 8 != 0 || seq <= 0 || rows % seq != 0) return QI_ERR_ARG;
  const int blocks = (rows + 7) / 8;
  if (head_dim == 128) {
    qi_launch(qknorm_rope_kernel<128>, dim3(blocks), dim3(256), 0, stream,
        (__nv_bfloat16*)qkv, rows, ld, heads, seq, s_txt, (const __nv_bfloat16*)wq_img,
        (const __nv_bfloat16*)wk_img, (const __nv_bfloat16*)wq_txt, (const __nv_bfloat16*)wk_txt,
        (const float2*)freqs_cos_sin, eps);
  } else if (head_dim == 64) {
    qi_launch(qknorm_rope_kernel<64>, dim3(blocks), dim3(256), 0, stream,
        (__nv_bfloat16*)qkv, rows, ld, heads, seq, s_txt, (const __nv_bfloat16*)wq_img,
        (const __nv_bfloat16*)wk_img, (const __nv_bfloat16*)wq_txt, (const __nv_bfloat16*)wk_txt,
        (const float2*)freqs_cos_sin, eps);
  } else {
    return QI_ERR_UNSUPPORTED;
  }
  QI_CHECK_LAUNCH();
  return QI_OK;
}

extern "C" int qi_qknorm_rope_scatter(const void* qkv, int rows, int ld, int heads, int head_dim,
                                      int seq, int s_txt, const void* wq_img, const void* wk_img,
                                      const void* wq_txt, const void* wk_txt,
                                      const void* freqs_cos_sin, float eps,
                                      const qi_peer_scatter_t* dst, cudaStream_t stream) {
  if (!qkv || !wq_img || !wk_img || !wq_txt || !wk_txt || !freqs_cos_sin || !dst) return QI_ERR_ARG;
  if (rows <= 0) return QI_OK;
  if (ld % 8 != 0 || seq <= 0 || rows % seq != 0) return QI_ERR_ARG;
  if (dst->n_peers < 1 || dst->n_peers > 8 || dst->my_index < 0 || dst->my_index >= dst->n_peers ||
      heads % dst->n_peers != 0)
    return QI_ERR_ARG;
  ScatterDst D;
  for (int p = 0; p < 8; ++p) {
    D.base[p] = p < dst->n_peers ? (__nv_bfloat16*)dst->base[p] : nullptr;
    if (p < dst->n_peers && (!dst->base[p] || (reinterpret_cast<uintptr_t>(dst->base[p]) & 15)))
      return QI_ERR_ARG;
  }
  D.n = dst->n_peers;
  D.heads_per_rank = heads / dst->n_peers;
  D.s_total = dst->row_prefix[dst->n_peers];
  D.row_offset = dst->row_prefix[dst->my_index];
  if (dst->row_prefix[dst->my_index + 1] - D.row_offset != seq) return QI_ERR_ARG;
  const int blocks = (rows + 7) / 8;
  if (head_dim == 128) {
    qi_launch(qknorm_rope_scatter_kernel<128>, dim3(blocks), dim3(256), 0, stream,
        (const __nv_bfloat16*)qkv, rows, ld, heads, seq, s_txt, (const __nv_bfloat16*)wq_img,
        (const __nv_bfloat16*)wk_img, (const __nv_bfloat16*)wq_txt, (const __nv_bfloat16*)wk_txt,
        (const float2*)freqs_cos_sin, eps, D);
  } else if (head_dim == 64) {
    qi_launch(qknorm_rope_scatter_kernel<64>, dim3(blocks), dim3(256), 0, stream,
        (const __nv_bfloat16*)qkv, rows, ld, heads, seq, s_txt, (const __nv_bfloat16*)wq_img,
        (const __nv_bfloat16*)wk_img, (const __nv_bfloat16*)wq_txt, (const __nv_bfloat16*)wk_txt,
        (const float2*)freqs_cos_sin, eps, D);
  } else {
    return QI_ERR_UNSUPPORTED;
  }
  QI_CHECK_LAUNCH();
  return QI_OK;
}

extern "C" int qi_cfg_euler_step(const void* cond, const void* uncond, const void* latents,
                                 void* out, long long tokens, int channels, float true_cfg_scale,
                                 const float* dt_dev, float dt, cudaStream_t stream) {
  if (!cond || !latents || !out) return QI_ERR_ARG;
  if (channels != 64) return QI_ERR_UNSUPPORTED;
  if (tokens <= 0) return QI_OK;
  const long long threads = tokens * 8;
  const int blocks = (int)((threads + 255) / 256);
  qi_launch(cfg_euler_kernel, dim3(blocks), dim3(256), 0, stream, (const __nv_bfloat16*)cond,
                                               (const __nv_bfloat16*)uncond,
                                               (const __nv_bfloat16*)latents, (__nv_bfloat16*)out,
                                               tokens, true_cfg_scale, dt_dev, dt);
  QI_CHECK_LAUNCH();
  return QI_OK;
}

extern "C" int qi_cfg_euler_step_f32(const float* cond, const float* uncond, const float* latents, float* out,
                                     void* out_bf16, long long tokens, int channels, float true_cfg_scale,
                                     const float* dt_dev, float dt, cudaStream_t stream) {
  if (!cond || !latents || !out) return QI_ERR_ARG;
  if (channels != 64) return QI_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(cond) | reinterpret_cast<uintptr_t>(uncond) | reinterpret_cast<uintptr_t>(latents) |
       reinterpret_cast<uintptr_t>(out)) & 15)
    return QI_ERR_ARG;
  if (tokens <= 0) return QI_OK;
  const long long threads = tokens * 16;
  const int blocks = (int)((threads + 255) / 256);
  qi_launch(cfg_euler_f32_kernel, dim3(blocks), dim3(256), 0, stream, cond, uncond, latents, out,
            (__nv_bfloat16*)out_bf16, tokens, true_cfg_scale, dt_dev, dt);
  QI_CHECK_LAUNCH();
  return QI_OK;
}

extern "C" int qi_small_linear(const void* x, const void* W, const void* bias, void* y, int M,
                               int N, int K, int ld_x, int ld_y, int silu_in, int silu_out,
                               cudaStream_t stream) {
  if (!x || !W || !y) return QI_ERR_ARG;
  if (M <= 0 || N <= 0) return QI_OK;
  if (K % 8 != 0 || ld_x % 8 != 0) return QI_ERR_ARG;
  dim3 grid((N + 7) / 8, (M + 3) / 4);
  qi_launch(small_linear_kernel<4, false>, grid, dim3(256), 0, stream, x, (const __nv_bfloat16*)W,
            (const __nv_bfloat16*)bias, y, M, N, K, ld_x, ld_y, silu_in, silu_out);
  QI_CHECK_LAUNCH();
  return QI_OK;
}

extern "C" int qi_small_linear_f32(const float* x, const void* W, const void* bias, float* y, int M, int N, int K,
                                   int ld_x, int ld_y, int silu_in, int silu_out, cudaStream_t stream) {
  if (!x || !W || !y) return QI_ERR_ARG;
  if (M <= 0 || N <= 0) return QI_OK;
  if (K % 8 != 0 || ld_x % 4 != 0 || (reinterpret_cast<uintptr_t>(x) & 15)) return QI_ERR_ARG;
  dim3 grid((N + 7) / 8, (M + 3) / 4);
  qi_launch(small_linear_kernel<4, true>, grid, dim3(256), 0, stream, (const void*)x, (const __nv_bfloat16*)W,
            (const __nv_bfloat16*)bias, (void*)y, M, N, K, ld_x, ld_y, silu_in, silu_out);
  QI_CHECK_LAUNCH();
  return QI_OK;
}

extern "C" int qi_timestep_sinusoid(const float* t, void* out, int M, cudaStream_t stream) {
  if (!t || !out) return QI_ERR_ARG;
  if (M <= 0) return QI_OK;
  qi_launch(timestep_sinusoid_kernel<false>, dim3((M * 128 + 127) / 128), dim3(128), 0, stream, t, out, M);
  QI_CHECK_LAUNCH();
  return QI_OK;
}

extern "C" int qi_timestep_sinusoid_f32(const float* t, float* out, int M, cudaStream_t stream) {
  if (!t || !out) return QI_ERR_ARG;
  if (M <= 0) return QI_OK;
  qi_launch(timestep_sinusoid_kernel<true>, dim3((M * 128 + 127) / 128), dim3(128), 0, stream, t, (void*)out, M);
  QI_CHECK_LAUNCH();
  return QI_OK;
}
