/* qi_b200.h — C ABI of the B200-native Qwen-Image denoising hot path (libqi_b200.so).
 *
 * The reference (AliciaCW/vllm-omni-draft) has no FFI of its own: its hot path is
 * `self.transformer(hidden_states=..., encoder_hidden_states=..., timestep=..., ...)`
 * (src/qwen_image/runner_adapter.py:128-137, src/qwen_image_gen/model.py:122-129) and
 * `pipe(**gen_kwargs)` (src/performance_enhancement/test_performance.py:277), both executed by the
 * un-vendored `diffusers` package on ATen/cuBLAS kernels.  These entry points are the operators
 * that call decomposes into (SURVEY.md §8b "What the C-ABI replacement must export"); each comment
 * names the upstream operator it replaces.  INTEGRATION.md shows the ctypes binding.
 *
 * Conventions: every pointer is a DEVICE pointer unless it says "host"; all activations/weights are
 * bf16 (uint16 storage), row-major, innermost dimension contiguous; `ld*` / `*_stride` are in
 * ELEMENTS; every function enqueues on `stream` and returns 0 or a negative qi_status.  No
 * allocation happens inside the library; the caller owns every buffer.  There is no CPU fallback:
 * on a machine without an sm_100 GPU every compute entry point returns QI_ERR_CUDA.
 */
#ifndef QI_B200_H_
#define QI_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* qi_stream_t; /* == cudaStream_t */

enum qi_status {
  QI_OK = 0,
  QI_ERR_ARG = -1,         /* malformed argument (null pointer, bad alignment/shape) */
  QI_ERR_CUDA = -2,        /* CUDA runtime error at launch (see qi_last_cuda_error) */
  QI_ERR_DRIVER = -3,      /* tensor-map encode failed */
  QI_ERR_UNSUPPORTED = -4  /* shape outside the compiled instantiations */
};

enum qi_epilogue {
  QI_EPI_BIAS = 0,           /* C = A W^T + b                         (to_q/k/v, add_*_proj, img_in, ...) */
  QI_EPI_BIAS_GELU_TANH = 1, /* C = gelu_tanh(A W^T + b)              (FeedForward net.0: GELU(approximate="tanh")) */
  QI_EPI_GATED_RESIDUAL = 2, /* C = R + gate[batch] * (A W^T + b)     (x = x + gate * attn/mlp output) */
  QI_EPI_QKV_NORM_ROPE = 3   /* C = [rope(rmsnorm(q)) | rope(rmsnorm(k)) | v] of A W_qkv^T + b: the fused to_q/k/v
                              * projection with norm_q/norm_k (RMSNorm over head_dim) and the MSRoPE rotation
                              * applied per head vector in the epilogue (QwenDoubleStreamAttnProcessor2_0) */
};

int qi_version(void);
const char* qi_last_cuda_error(void); /* host string, valid until the next failing call */
int qi_device_sm_count(void);
/* Programmatic dependent launch (every kernel of the library is launched with the programmatic
 * stream-serialization attribute and gates its global-memory work on griddepcontrol.wait, so a
 * kernel's prologue overlaps its predecessor's tail; CUDA-graph capture keeps the edges).  On by
 * default; 0 turns the attribute off (also: QI_PDL=0 in the environment). */
int qi_set_pdl(int on);

/* ---- K3: tcgen05 bf16 GEMM with fused epilogues (replaces nn.Linear + the elementwise ops after it).
 * Up to two independent problems (the img and txt streams of one MMDiT block) run in ONE persistent
 * launch.  A is [batch, m, k] (row stride lda, batch stride a_batch_stride), W is the nn.Linear
 * weight [n, k] contiguous, C is [batch, m, n] (ldc, c_batch_stride).  k % 64 == 0, n % 8 == 0.
 * A zero-initialised struct with a/w/c/m/batch/n/k set is a plain C = A W^T. */
struct qi_peer_scatter; /* defined with the Ulysses exchange entry points below */
struct qi_peer_signal;
typedef struct {
  const void* a;
  long long lda;
  long long a_batch_stride;
  const void* w;
  const void* bias; /* [n] or NULL */
  void* c;
  long long ldc;
  long long c_batch_stride;
  const void* residual; /* QI_EPI_GATED_RESIDUAL: R, may alias c */
  long long ld_res;
  long long res_batch_stride;
  const void* gate; /* QI_EPI_GATED_RESIDUAL: [batch, n] */
  long long gate_batch_stride;
  int m;     /* rows per batch item */
  int batch;
  int n;
  int k;
  int epilogue; /* enum qi_epilogue */
  /* Optional K-chunked A (Ulysses: attention output arrives as [head_group][m][k_chunk]):
   * element (row, kk) of chunk c lives at a + c*a_k_chunk_stride + row*lda + kk.  0 = A is plain [m, k]. */
  int a_k_chunk;
  long long a_k_chunk_stride;
  /* QI_EPI_QKV_NORM_ROPE only: n == 3 * heads * head_dim, columns [q | k | v], head_dim in {64, 128}. */
  const void* qk_norm_wq;    /* [head_dim] bf16 RMSNorm weight of the q heads */
  const void* qk_norm_wk;    /* [head_dim] bf16 RMSNorm weight of the k heads */
  const void* rope_cos_sin;  /* fp32 [>= rope_pos0 + m, head_dim/2, 2] (cos, sin); row r of every batch item uses
                              * table row rope_pos0 + r (joint order, text rows first) */
  int rope_pos0;
  int heads;
  int head_dim;
  float qk_norm_eps;
  /* Optional Ulysses scatter (NULL = store into c): head vector (b, row, j in {q,k,v}, head h) goes to
   * base[h / (heads/P)], laid out HEAD-MAJOR [batch, 3, heads/P, row_prefix[P], head_dim] (the layout
   * qi_qknorm_rope_scatter writes), at sequence row row_prefix[my_index] + rope_pos0 + row.  c may be NULL then.
   * scatter->notify (problem 0's) is signalled by the launch's last CTA. */
  const struct qi_peer_scatter* scatter;
  /* C (and the residual R of QI_EPI_GATED_RESIDUAL) are fp32 instead of bf16; ldc / ld_res / batch strides stay in
   * ELEMENTS.  QI_EPI_BIAS and QI_EPI_GATED_RESIDUAL only.  The MMDiT residual streams are kept in fp32 this way:
   * 60 blocks x 2 in-place bf16 roundings otherwise accumulate to a relative error of 1.9e-2. */
  int c_fp32;
  int gate_fp32; /* QI_EPI_GATED_RESIDUAL: gate is fp32 [batch, n] (a row of the fp32 modulation table) */
  /* Problem 0 only, host pointer or NULL: the A operands of this launch were stored by peers (the attention
   * scatter); the TMA producers poll these flags before their first load (see qi_peer_signal_t). */
  const struct qi_peer_signal* wait;
} qi_gemm_problem_t;

int qi_gemm_bf16(const qi_gemm_problem_t* problems /* host */, int num_problems /* 1 or 2 */,
                 qi_stream_t stream);
/* Tile width of the persistent GEMM: 0 = automatic (cheapest wave schedule over 148 SMs), or one of
 * 128 / 160 / 192 / 224 / 256 to pin it (tests and tile sweeps; QI_EPI_QKV_NORM_ROPE needs a multiple of head_dim).
 * The qi_gemm_set_* / qi_attention_set_variant functions write process-wide state that later launches read without
 * a lock: call them from the thread that launches, before the launches they are meant for (setup / tuning time). */
int qi_gemm_set_tile_n(int bn);
/* CTA-pair (tcgen05 cta_group::2, 256-row tiles on a 2-CTA cluster) vs single-CTA tiles:
 * -1 = automatic, 0 = never, 1 = always. */
int qi_gemm_set_cta_pair(int mode);
/* Measured schedule for problems of exactly these dims (m, batch, n, k, epilogue of each problem): later
 * qi_gemm_bf16 calls with matching dims use (pair, bn) instead of the wave-count heuristic.  The host
 * layer's autotuner (ops.tune_gemm) times every candidate once per new shape, before graph capture. */
int qi_gemm_set_schedule(const qi_gemm_problem_t* problems /* host */, int num_problems, int pair, int bn);
/* Same, with the tile rasterisation (see qi_gemm_set_group_m) as part of the measured schedule; -1 = default. */
int qi_gemm_set_schedule_ex(const qi_gemm_problem_t* problems /* host */, int num_problems, int pair, int bn,
                            int group_m);
int qi_gemm_clear_schedules(void);
/* Tile rasterisation of the persistent GEMM: supertiles of `group_m` row tiles x all column tiles (a wave of
 * concurrently running tiles is then a ~square patch of the tile grid); 0 = column-major order (row tile fastest),
 * -1 = default (env QI_GEMM_GROUP_M, else round(sqrt(workers))).  Tuning knob; single-threaded setup call like the
 * other qi_gemm_set_* functions (they write process-wide state read by later launches). */
int qi_gemm_set_group_m(int group_m);

/* ---- K4: joint text+image attention, non-causal, no mask, softmax scale 1/sqrt(head_dim)
 * (replaces dispatch_attention_fn / F.scaled_dot_product_attention in QwenDoubleStreamAttnProcessor2_0).
 * q, k, v, o are [batch, seq, heads, head_dim] views with a common row stride ld (elements between
 * consecutive tokens) and batch stride; head_dim in {64, 128}.  Queries 0..seq_q attend keys 0..seq_kv of the
 * k/v views (Ulysses passes the full-sequence K/V of its local heads); partial last tiles are masked in-kernel. */
int qi_attention_fwd(const void* q, const void* k, const void* v, void* o, int batch, int seq_q,
                     int seq_kv, int heads, int head_dim, long long ld_q, long long ld_kv,
                     long long ld_o, long long q_batch_stride, long long kv_batch_stride,
                     long long o_batch_stride, float softmax_scale, qi_stream_t stream);

/* ---- K1: LayerNorm(no affine, eps) fused with adaLN modulation: y = LN(x) * (1 + scale[b]) + shift[b]
 * (replaces img_norm1/2, txt_norm1/2 + QwenImageTransformerBlock._modulate, and AdaLayerNormContinuous).
 * shift/scale point at the first element of batch 0; batch b is mod_batch_stride elements further. */
int qi_adaln_modulate(const void* x, void* y, int rows, int d, int ld_x, int ld_y, const void* shift,
                      const void* scale, long long mod_batch_stride, int rows_per_batch, float eps,
                      qi_stream_t stream);

/* Same kernel over up to two row sets in ONE launch (the img and txt streams of a block). */
typedef struct {
  const void* x;
  void* y;
  int rows;
  int ld_x;
  int ld_y;
  const void* shift;
  const void* scale;
  long long mod_batch_stride;
  int rows_per_batch;
  int x_fp32; /* x rows are fp32 (the residual streams; ld_x still in elements); both row sets of a launch must agree */
  int mod_fp32; /* shift / scale are fp32 vectors (rows of the fp32 modulation table); both row sets must agree */
} qi_norm_rows_t;
int qi_adaln_modulate_multi(const qi_norm_rows_t* segs /* host */, int num_segs /* 1 or 2 */, int d, float eps,
                            qi_stream_t stream);

/* RMSNorm with learned weight (replaces txt_norm = RMSNorm(joint_attention_dim, eps=1e-6)). */
int qi_rmsnorm_affine(const void* x, void* y, int rows, int d, int ld_x, int ld_y, const void* weight,
                      float eps, qi_stream_t stream);

/* ---- K2: per-head RMSNorm of q and k (norm_q/norm_k for image rows, norm_added_q/k for text rows)
 * fused with the MSRoPE complex rotation (apply_rotary_emb_qwen, use_real=False), in place on the
 * joint qkv buffer [rows = batch*seq, ld] whose row is [q(heads*hd) | k(heads*hd) | v(...)].
 * freqs_cos_sin: fp32 [seq, head_dim/2, 2] = (cos, sin), joint order (text rows first). */
int qi_qknorm_rope(void* qkv, int rows, int ld, int heads, int head_dim, int seq, int s_txt,
                   const void* wq_img, const void* wk_img, const void* wq_txt, const void* wk_txt,
                   const void* freqs_cos_sin, float eps, qi_stream_t stream);

/* ---- Fused Ulysses exchange over peer memory (NVLink / NVSwitch; one process per GPU).
 * The reference has no parallelism; this is the head<->sequence all-to-all of
 * SURVEY §8e folded into the kernels on either side of it.  `base[p]` is rank p's buffer mapped
 * into this process (qi_peer_import, or any other peer mapping); rank p owns heads
 * [p*heads/P, (p+1)*heads/P) and sequence rows [row_prefix[p], row_prefix[p+1]) of the rank-major
 * joint sequence (ragged shards allowed). */
/* Flags of one exchange, folded into the kernels on either side of it (no qi_peer_barrier launch in between):
 * as `notify`, the LAST CTA of the scattering kernel bumps state[0] and release-increments slots[p][my_index] on
 * every peer once all of the kernel's peer stores are ordered before it; as `wait`, the consuming kernel's loader
 * thread acquire-polls slots[my_index][p] >= state[0] for every p before its first load of exchanged data (so the
 * consumer must follow, in stream order, the producer of the same exchange on the same rank, and every rank must
 * run the same sequence of exchanges).  A poll that exceeds timeout_ms sets state[2] and returns; once set, later
 * waits return immediately (the host checks it, PeerExchange.timed_out()). */
typedef struct qi_peer_signal {
  void* const* slots;   /* DEVICE array [n_peers]: slots[p] = rank p's uint32[n_peers] slot array (peer mapped) */
  unsigned int* state;  /* local DEVICE words, zero-initialised: [0] epoch, [1] CTA arrival counter, [2] timeout flag */
  int n_peers, my_index;
  unsigned int timeout_ms; /* 0: 2000 */
} qi_peer_signal_t;

typedef struct qi_peer_scatter {
  int n_peers;       /* P <= 8 */
  int my_index;      /* this rank's position in the group */
  void* base[8];
  int row_prefix[9]; /* row_prefix[P] = total sequence length */
  long long chunk_stride; /* attention scatter only: elements between head-group chunks at the destination */
  const qi_peer_signal_t* notify; /* host pointer or NULL: signal the peers when this kernel's stores are done */
  const qi_peer_signal_t* wait;   /* host pointer or NULL (qi_attention_fwd_scatter): flags of the exchange that
                                     delivered this kernel's q/k/v */
} qi_peer_scatter_t;

/* K2 with scatter: qkv is this rank's [batch*seq, ld] row block ([q|k|v], all heads); normalised and
 * rotated q/k vectors and untouched v vectors are stored to base[owner of the head], laid out HEAD-MAJOR
 * [batch, 3, heads/P, row_prefix[P], head_dim], at sequence row row_prefix[my_index] + s (one head's rows are
 * adjacent at the destination: peer writes of 128-byte segments reach 574 GB/s at a 256-byte row stride
 * against 324 GB/s at the 3*heads/P*head_dim*2-byte stride of a token-major buffer; tools/p2p_bw.py).
 * qi_gemm_bf16's QI_EPI_QKV_NORM_ROPE epilogue with `scatter` writes the same layout. */
int qi_qknorm_rope_scatter(const void* qkv, int rows, int ld, int heads, int head_dim, int seq, int s_txt,
                           const void* wq_img, const void* wk_img, const void* wq_txt, const void* wk_txt,
                           const void* freqs_cos_sin, float eps, const qi_peer_scatter_t* dst /* host */,
                           qi_stream_t stream);

/* K4 with scatter: attention over the full sequence for this rank's heads; q/k/v element (b, s, h, :) sits
 * at b*batch_stride + s*ld + h*head_stride (so head-major buffers are read in place); output row r (owned by
 * rank p) is stored to base[p] + my_index*chunk_stride + b*o_batch_stride + (r - row_prefix[p])*ld_o
 * + head*o_head_stride, i.e. into the K-chunked A operand of p's out-projection GEMM (with
 * ld_o = head_dim and o_head_stride = batch*S_max*head_dim every (source rank, head) is one K-chunk of
 * head_dim columns whose rows are contiguous). */
int qi_attention_fwd_scatter(const void* q, const void* k, const void* v, int batch, int seq_q, int seq_kv,
                             int heads, int head_dim, long long ld_q, long long ld_kv, long long ld_o,
                             long long q_batch_stride, long long kv_batch_stride, long long o_batch_stride,
                             long long q_head_stride, long long kv_head_stride, long long o_head_stride,
                             float softmax_scale, const qi_peer_scatter_t* dst /* host */, qi_stream_t stream);
/* Softmax form of the attention kernel, one A/B switch: -1 = default (env QI_ATTN_VARIANT, else 11), otherwise
 * EMU + 8 * SPLIT with EMU in {0, 3} = how many of every 8 element pairs evaluate 2^x with the degree-3 FMA-pipe
 * polynomial instead of MUFU.EX2 and SPLIT in {0, 1} = P handed to the P*V MMA in two halves; i.e. 0, 3, 8 or 11.
 * Anything else is QI_ERR_ARG.  All four agree with fp32 SDPA to the same tolerance (tests/test_ops_gpu.py); 11 is
 * the fastest on the benchmarked shapes.  The other kernels and mixes that were built and measured slower
 * (profiles/) are not part of this library: tools/experiments/Makefile builds them into a
 * separate libqi_b200_exp.so for the sweep tools. */
int qi_attention_set_variant(int variant);

/* Tuning aid (measurement only; not needed by a drop-in caller): bit 0 skips the fp32 epilogue's residual loads,
 * bit 1 its result stores, bit 2 its bias/gate loads (wrong results; attributes epilogue time, tools/ab_gemm.py). */
int qi_gemm_set_debug(int mask);


/* Peer memory: device allocations that can be mapped by the other ranks of the node (CUDA IPC).
 * handle64 is a 64-byte host blob to ship to the peers (e.g. torch.distributed.all_gather_object). */
int qi_peer_alloc(void** ptr /* host out */, unsigned long long bytes); /* zero-filled */
int qi_peer_free(void* ptr);
int qi_peer_export(void* ptr, void* handle64 /* host */);
int qi_peer_import(const void* handle64 /* host */, void** ptr /* host out */);
int qi_peer_unmap(void* ptr);
/* Flag barrier across the ranks of a group, enqueued on `stream` (graph-capturable): orders every
 * peer store issued by earlier kernels of the stream before the flag, then waits for all peers.
 * signal_ptrs_dev: DEVICE array of n_peers pointers, entry p = rank p's u32[n_peers] signal array
 * (zero-initialised, peer-mapped).  state_dev: this rank's u32[2] {epoch, sticky timeout flag}. */
int qi_peer_barrier(void* const* signal_ptrs_dev, int n_peers, int my_index, uint32_t* state_dev,
                    unsigned int timeout_ms, qi_stream_t stream);

/* qi_peer_barrier without any device memory of the caller's for its bookkeeping.  This entry point has no
 * counterpart in the reference; it exists for hosts that cannot allocate on the GPU when they need a barrier:
 * a serving worker that gave all remaining HBM to the engine's block pool after memory profiling (a late
 * cudaMalloc fails or eats into that pool), a denoising loop that cannot afford the device-wide synchronisation
 * cudaMalloc / cudaFree imply, or a C / Go / Rust host that binds this library and owns no device allocator.
 * The pointer table and the {epoch, timeout flag} words that qi_peer_barrier takes as signal_ptrs_dev / state_dev
 * live in a fixed 256-byte array inside the library instead (present on every GPU that loads the library, until
 * the process exits).  The call itself performs no cudaMalloc / cudaFree and calls nothing that allocates.
 *   signal_ptrs   HOST array of n_peers DEVICE pointers, entry p = rank p's u32[n_peers] signal array
 *                 (zero-initialised, peer-mapped — still the caller's: they are the barrier's input/output).
 *                 Read before the call returns.
 *   n_peers       1 .. QI_PEER_BARRIER_STATIC_MAX_PEERS; more is refused with QI_ERR_ARG and a message in
 *                 qi_last_cuda_error() (qi_peer_barrier has the same limit today; here it is the table's size).
 *   restart       The array is shared by every call on the device and keeps what the last call left in it.  The
 *                 epoch must carry over from one barrier to the next on the same signal arrays, so it is NOT
 *                 cleared by default.  Pass non-zero on the first barrier over freshly zeroed signal arrays
 *                 (every rank, same call): epoch and timeout flag are then reset on `stream` first.
 *   state_out_dev DEVICE u32[2] or NULL: receives {epoch, timeout flag} after this barrier (what qi_peer_barrier
 *                 leaves in state_dev), so a timeout can be seen by the host.
 * One group per device at a time: two barriers that use this entry point and overlap on the same GPU (different
 * streams, or two groups) race on the shared words.  The first call on each device looks up the array's device
 * address (cudaGetSymbolAddress, with that device current); make that first call outside any stream capture. */
#define QI_PEER_BARRIER_STATIC_MAX_PEERS 8
int qi_peer_barrier_static(void* const* signal_ptrs /* host */, int n_peers, int my_index, int restart,
                           uint32_t* state_out_dev, unsigned int timeout_ms, qi_stream_t stream);

/* ---- K5: true-CFG combine + per-token norm rescale + FlowMatch Euler update in one pass
 * (replaces QwenImagePipeline's comb_pred/cond_norm/noise_norm lines and
 * FlowMatchEulerDiscreteScheduler.step).  uncond == NULL skips CFG.  dt = sigma_next - sigma is read
 * from dt_dev (device fp32) when non-NULL, else from the scalar.  channels must be 64. */
int qi_cfg_euler_step(const void* cond, const void* uncond, const void* latents, void* out,
                      long long tokens, int channels, float true_cfg_scale, const float* dt_dev,
                      float dt, qi_stream_t stream);

/* Same step with fp32 predictions, fp32 latents and fp32 arithmetic throughout (what the pipeline uses: true CFG
 * amplifies the rounding of the two predictions by 2s-1, and a bf16 latent re-rounded every step drifts from the fp32
 * trajectory like sqrt(steps) * 1.7e-3).  out: the fp32 master copy of the new latents; out_bf16 (may be NULL): the
 * bf16 copy the next forward reads. */
int qi_cfg_euler_step_f32(const float* cond, const float* uncond, const float* latents, float* out, void* out_bf16,
                          long long tokens, int channels, float true_cfg_scale, const float* dt_dev, float dt,
                          qi_stream_t stream);

/* ---- small-M linear (M <= a few rows): y = act_out(act_in(x) W^T + b); used for the timestep
 * embedder, the per-block modulation Linear(SiLU(temb)) and norm_out.linear.  HBM-bound on W. */
int qi_small_linear(const void* x, const void* w, const void* bias, void* y, int m, int n, int k,
                    int ld_x, int ld_y, int silu_in, int silu_out, qi_stream_t stream);

/* Same with fp32 activations in and out and no intermediate rounding (weights stay bf16): the timestep embedder
 * and the modulation Linear feed every block of every step, so their rounding error is coherent across the whole
 * forward; the fp32 form is what the transformer uses by default. */
int qi_small_linear_f32(const float* x, const void* w, const void* bias, float* y, int m, int n, int k,
                        int ld_x, int ld_y, int silu_in, int silu_out, qi_stream_t stream);

/* Timesteps(256, flip_sin_to_cos=True, downscale_freq_shift=0, scale=1000): t fp32 [m] -> bf16 [m,256]. */
int qi_timestep_sinusoid(const float* t, void* out, int m, qi_stream_t stream);
int qi_timestep_sinusoid_f32(const float* t, float* out, int m, qi_stream_t stream); /* fp32 [m,256] */

/* ---- host-side integer tables (bit-exact parts of the path; no GPU needed).
 * MSRoPE position ids (QwenEmbedRope): img_pos [sum F*H*W, 3], txt_pos [txt_len, 3] (int32, host).
 * shapes: num_images triples (F,H,W).  Returns 0, or QI_ERR_ARG if a text position would
 * exceed the 4096-entry table. */
int qi_rope_position_ids(const int* shapes /* host */, int num_images, int txt_len,
                         int32_t* img_pos /* host */, int32_t* txt_pos /* host */);
/* Token permutation of _pack_latents: for packed element index i of [ (Hl/2)(Wl/2), C*4 ] gives the
 * flat source index into [C, Hl, Wl].  idx: host int64 [(Hl/2)*(Wl/2)*C*4]. */
int qi_pack_latents_index(int channels, int hl, int wl, int64_t* idx /* host */);

#ifdef __cplusplus
}
#endif
#endif /* QI_B200_H_ */
