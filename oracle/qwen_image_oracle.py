"""ORACLE — test infrastructure only.  PARITY UNPINNED.

Plain-PyTorch restatement of the Qwen-Image denoising hot path that the reference
(AliciaCW/vllm-omni-draft) *calls* but does not *contain*: the arithmetic lives in the
un-vendored, un-pinned third-party package ``diffusers`` (>= 0.35 implied by
``set_attention_backend("_flash_3_hub")``, reference src/performance_enhancement/
test_performance.py:320-325).  ``diffusers`` is not installable here (no wheel, no network) and
the reference's own tests hold no golden tensors for this path, so this
restatement is pinned by neither golden vectors nor a run of the real thing: **parity
unpinned**.  It follows the published algorithm of diffusers 0.35's
``transformer_qwenimage.py`` / ``scheduling_flow_match_euler_discrete.py`` /
``pipeline_qwenimage.py`` and is anchored on the reference's call sites:

* transformer forward kwargs: src/qwen_image/runner_adapter.py:128-137,
  src/qwen_image_gen/model.py:122-129;
* pipeline call kwargs: src/performance_enhancement/test_pipeline.py:247-266,
  test_performance.py:244-266, 277;
* operator inventory: src/doc/archive/QwenImage_Diffusers_Interface.md:7-21, 213-220.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this file.  The product package never does.

Everything is written op-by-op the way the upstream modules execute it, so that running it in
bf16 reproduces upstream's rounding order (each eager op rounds to bf16; RMSNorm / RoPE /
scheduler.step upcast to fp32 exactly where upstream does).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------
# Config (Appendix A.0)
# --------------------------------------------------------------------------------------
@dataclass(frozen=True)
class OracleConfig:
    num_layers: int = 60
    attention_head_dim: int = 128
    num_attention_heads: int = 24
    joint_attention_dim: int = 3584
    in_channels: int = 64
    out_channels: int = 16
    patch_size: int = 2
    axes_dims_rope: Tuple[int, int, int] = (16, 56, 56)
    rope_theta: float = 10000.0
    guidance_embeds: bool = False

    @property
    def inner_dim(self) -> int:
        return self.num_attention_heads * self.attention_head_dim

    @property
    def proj_out_dim(self) -> int:
        return self.patch_size * self.patch_size * self.out_channels


def tiny_config() -> OracleConfig:
    """BASELINE.json configs[0]: 2 blocks, hidden 256, 4 heads (Dh 64)."""
    return OracleConfig(num_layers=2, attention_head_dim=64, num_attention_heads=4,
                        joint_attention_dim=128, axes_dims_rope=(8, 28, 28))


def qwen_image_20b_config() -> OracleConfig:
    return OracleConfig()


# --------------------------------------------------------------------------------------
# Random-init weights with diffusers parameter names (Appendix A.7)
# --------------------------------------------------------------------------------------
def init_state_dict(cfg: OracleConfig, seed: int = 0, dtype=torch.float32,
                    weight_std: float = 0.02, bias_std: float = 0.0,
                    norm_weight_std: float = 0.0) -> Dict[str, torch.Tensor]:
    """Linear.weight ~ N(0, 0.02^2), bias = 0, norm weights = 1 (fp32, then cast).
    Tests pass bias_std / norm_weight_std > 0 so that bias and affine paths are exercised."""
    g = torch.Generator().manual_seed(seed)
    d, dt, hd = cfg.inner_dim, cfg.joint_attention_dim, cfg.attention_head_dim
    sd: Dict[str, torch.Tensor] = {}

    def lin(name, out_f, in_f):
        sd[name + ".weight"] = (torch.randn(out_f, in_f, generator=g) * weight_std).to(dtype)
        b = torch.randn(out_f, generator=g) * bias_std if bias_std > 0 else torch.zeros(out_f)
        sd[name + ".bias"] = b.to(dtype)

    def normw(name, n):
        w = torch.ones(n)
        if norm_weight_std > 0:
            w = w + torch.randn(n, generator=g) * norm_weight_std
        sd[name + ".weight"] = w.to(dtype)

    lin("img_in", d, cfg.in_channels)
    normw("txt_norm", dt)
    lin("txt_in", d, dt)
    lin("time_text_embed.timestep_embedder.linear_1", d, 256)
    lin("time_text_embed.timestep_embedder.linear_2", d, d)
    for i in range(cfg.num_layers):
        p = f"transformer_blocks.{i}."
        lin(p + "img_mod.1", 6 * d, d)
        lin(p + "txt_mod.1", 6 * d, d)
        for n in ("to_q", "to_k", "to_v", "add_q_proj", "add_k_proj", "add_v_proj"):
            lin(p + "attn." + n, d, d)
        for n in ("norm_q", "norm_k", "norm_added_q", "norm_added_k"):
            normw(p + "attn." + n, hd)
        lin(p + "attn.to_out.0", d, d)
        lin(p + "attn.to_add_out", d, d)
        for s in ("img_mlp", "txt_mlp"):
            lin(p + s + ".net.0.proj", 4 * d, d)
            lin(p + s + ".net.2", d, 4 * d)
    lin("norm_out.linear", 2 * d, d)
    lin("proj_out", cfg.proj_out_dim, d)
    return sd


class ParamView:
    """State dict for the oracle that READS THE MODEL UNDER TEST: diffusers-named tensors fetched from any object
    with `diffusers_param(name)` (the CUDA model's fused parameters) and upcast on access.  Lets the oracle check a
    20B-parameter model on the GPU on the very same (bf16-rounded) weights without a second copy of them."""

    def __init__(self, model, dtype=torch.float32):
        self.model, self.dtype = model, dtype

    def __getitem__(self, name: str) -> torch.Tensor:
        return self.model.diffusers_param(name).to(self.dtype)


# --------------------------------------------------------------------------------------
# MSRoPE (Appendix A.5).  Integer position ids are the bit-exact part.
# --------------------------------------------------------------------------------------
ROPE_TABLE_LEN = 4096


def rope_position_ids(img_shapes: Sequence[Tuple[int, int, int]], max_txt_len: int,
                      scale_rope: bool = True) -> Tuple[torch.Tensor, torch.Tensor]:
    """Integer (frame, h, w) position of every image token and of every text token.

    Returns (img_pos [S_img, 3] int64, txt_pos [S_txt, 3] int64).  Mirrors
    QwenEmbedRope.forward/_compute_video_freqs: frame idx..idx+F-1, centred h/w ranges,
    text positions start at max(H//2, W//2) over all images, same value on all three axes."""
    rows = []
    max_vid_index = 0
    for idx, (f, h, w) in enumerate(img_shapes):
        fr = torch.arange(idx, idx + f, dtype=torch.int64)
        if scale_rope:
            hh = torch.cat([torch.arange(-(h - h // 2), 0), torch.arange(0, h // 2)]).to(torch.int64)
            ww = torch.cat([torch.arange(-(w - w // 2), 0), torch.arange(0, w // 2)]).to(torch.int64)
            max_vid_index = max(h // 2, w // 2, max_vid_index)
        else:
            hh = torch.arange(h, dtype=torch.int64)
            ww = torch.arange(w, dtype=torch.int64)
            max_vid_index = max(h, w, max_vid_index)
        grid = torch.stack([
            fr.view(f, 1, 1).expand(f, h, w),
            hh.view(1, h, 1).expand(f, h, w),
            ww.view(1, 1, w).expand(f, h, w)], dim=-1).reshape(f * h * w, 3)
        rows.append(grid)
    img_pos = torch.cat(rows, dim=0)
    t = torch.arange(max_vid_index, max_vid_index + max_txt_len, dtype=torch.int64)
    if max_vid_index + max_txt_len > ROPE_TABLE_LEN:
        raise ValueError("text positions exceed the 4096-entry RoPE table")
    txt_pos = torch.stack([t, t, t], dim=-1)
    return img_pos, txt_pos


def normalize_img_shapes(img_shapes):
    """QwenEmbedRope.forward: a list is indexed [0] (only the first sample's shapes are used);
    a bare (F,H,W) tuple is wrapped into a one-element list."""
    if isinstance(img_shapes, list):
        img_shapes = img_shapes[0]
    if not isinstance(img_shapes, list):
        img_shapes = [img_shapes]
    return [tuple(int(v) for v in s) for s in img_shapes]


def _rope_params(index: torch.Tensor, dim: int, theta: float) -> torch.Tensor:
    freqs = torch.outer(index, 1.0 / torch.pow(theta, torch.arange(0, dim, 2).to(torch.float32).div(dim)))
    return torch.polar(torch.ones_like(freqs), freqs)


def rope_freqs(cfg: OracleConfig, img_shapes, max_txt_len: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """complex64 rotation tables: (vid_freqs [S_img, Dh/2], txt_freqs [S_txt, Dh/2])."""
    img_pos, txt_pos = rope_position_ids(img_shapes, max_txt_len)

    def gather(pos):
        parts = []
        for a, dim in enumerate(cfg.axes_dims_rope):
            # upstream builds pos table over arange(4096) and neg table over [-4096..-1]; indexing a
            # row of either is the same as evaluating the outer product at that signed integer.
            parts.append(_rope_params(pos[:, a], dim, cfg.rope_theta))
        return torch.cat(parts, dim=1)

    return gather(img_pos), gather(txt_pos)


def apply_rope(x: torch.Tensor, freqs: torch.Tensor) -> torch.Tensor:
    """x [B,S,H,Dh]; adjacent (even, odd) pairs as complex; fp32 math; cast back."""
    xc = torch.view_as_complex(x.float().reshape(*x.shape[:-1], -1, 2))
    out = torch.view_as_real(xc * freqs.unsqueeze(1)).flatten(3)
    return out.type_as(x)


# --------------------------------------------------------------------------------------
# Building blocks
# --------------------------------------------------------------------------------------
def rms_norm(x: torch.Tensor, weight: Optional[torch.Tensor], eps: float = 1e-6) -> torch.Tensor:
    """diffusers RMSNorm: variance in fp32, x*rsqrt (fp32), cast to weight dtype, * weight."""
    in_dtype = x.dtype
    var = x.to(torch.float32).pow(2).mean(-1, keepdim=True)
    x = x * torch.rsqrt(var + eps)
    if weight is not None:
        if weight.dtype in (torch.float16, torch.bfloat16):
            x = x.to(weight.dtype)
        x = x * weight
    else:
        x = x.to(in_dtype)
    return x


def layer_norm(x: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    return F.layer_norm(x, (x.shape[-1],), None, None, eps)


def timestep_sinusoid(timestep: torch.Tensor) -> torch.Tensor:
    """Timesteps(256, flip_sin_to_cos=True, downscale_freq_shift=0, scale=1000) (A.2)."""
    half = 128
    exponent = -math.log(10000) * torch.arange(0, half, dtype=torch.float32, device=timestep.device) / half
    emb = timestep[:, None].float() * torch.exp(exponent)[None, :]
    emb = 1000 * emb
    return torch.cat([torch.cos(emb), torch.sin(emb)], dim=-1)


def time_embed(sd, timestep: torch.Tensor, dtype) -> torch.Tensor:
    proj = timestep_sinusoid(timestep).to(dtype)
    p = "time_text_embed.timestep_embedder."
    h = F.linear(proj, sd[p + "linear_1.weight"], sd[p + "linear_1.bias"])
    h = F.silu(h)
    return F.linear(h, sd[p + "linear_2.weight"], sd[p + "linear_2.bias"])


def _modulate(x, mod):
    shift, scale, gate = mod.chunk(3, dim=-1)
    return x * (1 + scale.unsqueeze(1)) + shift.unsqueeze(1), gate.unsqueeze(1)


def joint_attention(sd, p, cfg: OracleConfig, img, txt, img_freqs, txt_freqs):
    H, Dh = cfg.num_attention_heads, cfg.attention_head_dim
    a = p + "attn."

    def proj(x, n):
        return F.linear(x, sd[a + n + ".weight"], sd[a + n + ".bias"]).unflatten(-1, (H, Dh))

    iq, ik, iv = proj(img, "to_q"), proj(img, "to_k"), proj(img, "to_v")
    tq, tk, tv = proj(txt, "add_q_proj"), proj(txt, "add_k_proj"), proj(txt, "add_v_proj")
    iq, ik = rms_norm(iq, sd[a + "norm_q.weight"]), rms_norm(ik, sd[a + "norm_k.weight"])
    tq, tk = rms_norm(tq, sd[a + "norm_added_q.weight"]), rms_norm(tk, sd[a + "norm_added_k.weight"])
    iq, ik = apply_rope(iq, img_freqs), apply_rope(ik, img_freqs)
    tq, tk = apply_rope(tq, txt_freqs), apply_rope(tk, txt_freqs)
    q = torch.cat([tq, iq], dim=1).transpose(1, 2)      # text first
    k = torch.cat([tk, ik], dim=1).transpose(1, 2)
    v = torch.cat([tv, iv], dim=1).transpose(1, 2)
    o = F.scaled_dot_product_attention(q, k, v, attn_mask=None, dropout_p=0.0, is_causal=False)
    o = o.transpose(1, 2).flatten(2, 3).to(q.dtype)
    s_txt = txt.shape[1]
    t_out, i_out = o[:, :s_txt], o[:, s_txt:]
    i_out = F.linear(i_out, sd[a + "to_out.0.weight"], sd[a + "to_out.0.bias"])
    t_out = F.linear(t_out, sd[a + "to_add_out.weight"], sd[a + "to_add_out.bias"])
    return i_out, t_out


def mlp(sd, p, x):
    h = F.linear(x, sd[p + ".net.0.proj.weight"], sd[p + ".net.0.proj.bias"])
    h = F.gelu(h, approximate="tanh")
    return F.linear(h, sd[p + ".net.2.weight"], sd[p + ".net.2.bias"])


def block_forward(sd, i: int, cfg: OracleConfig, h, e, temb, img_freqs, txt_freqs):
    """QwenImageTransformerBlock.forward (A.3).  Returns (e, h)."""
    p = f"transformer_blocks.{i}."
    img_mod = F.linear(F.silu(temb), sd[p + "img_mod.1.weight"], sd[p + "img_mod.1.bias"])
    txt_mod = F.linear(F.silu(temb), sd[p + "txt_mod.1.weight"], sd[p + "txt_mod.1.bias"])
    img_mod1, img_mod2 = img_mod.chunk(2, dim=-1)
    txt_mod1, txt_mod2 = txt_mod.chunk(2, dim=-1)
    img_m, img_g1 = _modulate(layer_norm(h), img_mod1)
    txt_m, txt_g1 = _modulate(layer_norm(e), txt_mod1)
    a_img, a_txt = joint_attention(sd, p, cfg, img_m, txt_m, img_freqs, txt_freqs)
    h = h + img_g1 * a_img
    e = e + txt_g1 * a_txt
    img_m2, img_g2 = _modulate(layer_norm(h), img_mod2)
    h = h + img_g2 * mlp(sd, p + "img_mlp", img_m2)
    txt_m2, txt_g2 = _modulate(layer_norm(e), txt_mod2)
    e = e + txt_g2 * mlp(sd, p + "txt_mlp", txt_m2)
    return e, h


def transformer_forward(sd, cfg: OracleConfig, hidden_states, encoder_hidden_states,
                        encoder_hidden_states_mask, timestep, img_shapes=None, txt_seq_lens=None,
                        guidance=None, num_layers: Optional[int] = None,
                        return_intermediates: bool = False):
    """QwenImageTransformer2DModel.forward (A.1).  timestep is the pipeline's t/1000.
    The mask is threaded but not applied (diffusers 0.35 behaviour, attn_mask=None)."""
    B, S_img, _ = hidden_states.shape
    S_txt = encoder_hidden_states.shape[1]
    if img_shapes is None:
        side = int(round(math.sqrt(S_img)))
        if side * side != S_img:
            raise ValueError("img_shapes required for non-square token grids")
        img_shapes = [(1, side, side)]
    img_shapes = normalize_img_shapes(img_shapes)
    max_txt = max(txt_seq_lens) if txt_seq_lens is not None else S_txt
    dtype = hidden_states.dtype
    h = F.linear(hidden_states, sd["img_in.weight"], sd["img_in.bias"])
    timestep = timestep.to(dtype)
    e = rms_norm(encoder_hidden_states, sd["txt_norm.weight"])
    e = F.linear(e, sd["txt_in.weight"], sd["txt_in.bias"])
    temb = time_embed(sd, timestep, dtype)
    img_freqs, txt_freqs = rope_freqs(cfg, img_shapes, max_txt)
    img_freqs, txt_freqs = img_freqs.to(h.device), txt_freqs.to(h.device)
    inter = []
    L = cfg.num_layers if num_layers is None else num_layers
    for i in range(L):
        e, h = block_forward(sd, i, cfg, h, e, temb, img_freqs, txt_freqs)
        if return_intermediates:
            inter.append((h.clone(), e.clone()))
    emb = F.linear(F.silu(temb).to(dtype), sd["norm_out.linear.weight"], sd["norm_out.linear.bias"])
    scale, shift = torch.chunk(emb, 2, dim=1)        # (scale, shift): order differs from the blocks
    h = layer_norm(h) * (1 + scale)[:, None, :] + shift[:, None, :]
    out = F.linear(h, sd["proj_out.weight"], sd["proj_out.bias"])
    if return_intermediates:
        return out, inter, temb
    return out


# --------------------------------------------------------------------------------------
# Pipeline glue + FlowMatch Euler scheduler + true CFG (Appendix A.6)
# --------------------------------------------------------------------------------------
def pack_latents(latents: torch.Tensor) -> torch.Tensor:
    """[B,1,16,Hl,Wl] or [B,16,Hl,Wl] -> [B,(Hl/2)(Wl/2),64]  (pure index permutation)."""
    if latents.dim() == 5:
        latents = latents[:, 0] if latents.shape[1] == 1 else latents.squeeze(2)
    B, C, Hl, Wl = latents.shape
    x = latents.view(B, C, Hl // 2, 2, Wl // 2, 2).permute(0, 2, 4, 1, 3, 5)
    return x.reshape(B, (Hl // 2) * (Wl // 2), C * 4)


def unpack_latents(tokens: torch.Tensor, Hl: int, Wl: int) -> torch.Tensor:
    """Inverse of pack_latents -> [B,16,1,Hl,Wl]."""
    B, _, ch = tokens.shape
    x = tokens.view(B, Hl // 2, Wl // 2, ch // 4, 2, 2).permute(0, 3, 1, 4, 2, 5)
    return x.reshape(B, ch // 4, 1, Hl, Wl)


def latent_hw(height_px: int, width_px: int, vae_scale_factor: int = 8) -> Tuple[int, int]:
    return 2 * (int(height_px) // (vae_scale_factor * 2)), 2 * (int(width_px) // (vae_scale_factor * 2))


def calculate_shift(image_seq_len, base_seq_len=256, max_seq_len=8192, base_shift=0.5, max_shift=0.9):
    m = (max_shift - base_shift) / (max_seq_len - base_seq_len)
    b = base_shift - m * base_seq_len
    return image_seq_len * m + b


def flowmatch_sigmas(num_inference_steps: int, image_seq_len: int, shift_terminal: float = 0.02, sigmas=None):
    """FlowMatchEulerDiscreteScheduler.set_timesteps(sigmas=linspace(1,1/N,N), mu=...) with the
    Qwen-Image scheduler config (dynamic exponential shift + terminal stretch), numpy float32.
    Returns (timesteps [N] fp32, sigmas [N+1] fp32 with trailing 0)."""
    if sigmas is None:
        sigmas = np.linspace(1.0, 1.0 / num_inference_steps, num_inference_steps)
    sigmas = np.array(sigmas).astype(np.float32)      # pipeline's default, or the caller's `sigmas=` list
    mu = calculate_shift(image_seq_len)
    sigmas = math.exp(mu) / (math.exp(mu) + (1 / sigmas - 1) ** 1.0)
    one_minus = 1 - sigmas
    scale = one_minus[-1] / (1 - shift_terminal)
    sigmas = 1 - one_minus / scale
    sigmas = torch.from_numpy(np.asarray(sigmas)).to(torch.float32)
    timesteps = sigmas * 1000.0
    sigmas = torch.cat([sigmas, torch.zeros(1)])
    return timesteps, sigmas


def scheduler_step(model_output: torch.Tensor, sample: torch.Tensor, sigma: torch.Tensor,
                   sigma_next: torch.Tensor) -> torch.Tensor:
    """FlowMatchEulerDiscreteScheduler.step: upcast sample, x += dt*v, cast to v's dtype.
    `dt * model_output` is a 0-dim fp32 tensor times a bf16 tensor => evaluated in bf16."""
    sample = sample.to(torch.float32)
    dt = sigma_next - sigma
    prev = sample + dt * model_output
    return prev.to(model_output.dtype)


def true_cfg_combine(cond: torch.Tensor, uncond: torch.Tensor, scale: float) -> torch.Tensor:
    comb = uncond + scale * (cond - uncond)
    cond_norm = torch.norm(cond, dim=-1, keepdim=True)
    noise_norm = torch.norm(comb, dim=-1, keepdim=True)
    return comb * (cond_norm / noise_norm)


def denoise_loop(sd, cfg: OracleConfig, latents, prompt_embeds, prompt_mask,
                 negative_prompt_embeds=None, negative_prompt_mask=None,
                 num_inference_steps: int = 50, true_cfg_scale: float = 4.0,
                 img_shapes=None, return_all: bool = False):
    """QwenImagePipeline.__call__ denoising loop on packed latents [B,S_img,64]."""
    B, S_img, _ = latents.shape
    timesteps, sigmas = flowmatch_sigmas(num_inference_steps, S_img)
    do_cfg = true_cfg_scale > 1 and negative_prompt_embeds is not None
    txt_lens = prompt_mask.sum(dim=1).tolist()
    neg_lens = negative_prompt_mask.sum(dim=1).tolist() if do_cfg else None
    trace = []
    for i, t in enumerate(timesteps):
        timestep = t.expand(B).to(latents.dtype)
        c = transformer_forward(sd, cfg, latents, prompt_embeds, prompt_mask, timestep / 1000,
                                img_shapes=img_shapes, txt_seq_lens=txt_lens)
        if do_cfg:
            u = transformer_forward(sd, cfg, latents, negative_prompt_embeds, negative_prompt_mask,
                                    timestep / 1000, img_shapes=img_shapes, txt_seq_lens=neg_lens)
            v = true_cfg_combine(c, u, true_cfg_scale)
        else:
            v = c
        latents = scheduler_step(v, latents, sigmas[i], sigmas[i + 1])
        if return_all:
            trace.append(latents.clone())
    return (latents, trace) if return_all else latents


def edit_denoise_loop(sd, cfg: OracleConfig, latents, image_latents, prompt_embeds, prompt_mask,
                      negative_prompt_embeds=None, negative_prompt_mask=None,
                      num_inference_steps: int = 50, true_cfg_scale: float = 4.0,
                      img_shapes=None, return_all: bool = False):
    """QwenImageEditPipeline.__call__ denoising loop (what the reference's
    test_performance.py:229-307 / test_pipeline.py:123-124 actually drive): the packed reference-image
    latents ride behind the noise tokens on the sequence axis, ``img_shapes = [[(1,h,w), (1,h',w')]]``
    (RoPE frame index = image index), and only the first S_noise rows of the prediction are kept.
    mu / the sigma schedule use the noise sequence length."""
    B, S_noise, _ = latents.shape
    timesteps, sigmas = flowmatch_sigmas(num_inference_steps, S_noise)
    do_cfg = true_cfg_scale > 1 and negative_prompt_embeds is not None
    txt_lens = prompt_mask.sum(dim=1).tolist()
    neg_lens = negative_prompt_mask.sum(dim=1).tolist() if do_cfg else None
    trace = []
    for i, t in enumerate(timesteps):
        timestep = t.expand(B).to(latents.dtype)
        x = torch.cat([latents, image_latents], dim=1)
        c = transformer_forward(sd, cfg, x, prompt_embeds, prompt_mask, timestep / 1000,
                                img_shapes=img_shapes, txt_seq_lens=txt_lens)[:, :S_noise]
        if do_cfg:
            u = transformer_forward(sd, cfg, x, negative_prompt_embeds, negative_prompt_mask,
                                    timestep / 1000, img_shapes=img_shapes, txt_seq_lens=neg_lens)[:, :S_noise]
            v = true_cfg_combine(c, u, true_cfg_scale)
        else:
            v = c
        latents = scheduler_step(v, latents, sigmas[i], sigmas[i + 1])
        if return_all:
            trace.append(latents.clone())
    return (latents, trace) if return_all else latents


# --------------------------------------------------------------------------------------
# The reference's OWN placeholder loops (these lines are in /root/reference, not upstream)
# --------------------------------------------------------------------------------------
def reference_adapter_loop(sd, cfg: OracleConfig, latents, prompt_embeds, prompt_mask, num_steps: int,
                           img_shapes=None, txt_seq_lens=None, step_size: float = 1.0):
    """src/qwen_image/runner_adapter.py:64-68, 127-141: int64 timesteps linspace(1,0,N)*1000 passed
    straight to the transformer, ``latents - residual``.  latents [B,64,H,W] (or 5-D)."""
    timesteps = (torch.linspace(1.0, 0.0, steps=max(int(num_steps), 1)) * 1000).to(torch.long)
    return reference_gen_loop(sd, cfg, latents, prompt_embeds, prompt_mask, timesteps, step_size=step_size,
                              img_shapes=img_shapes, txt_seq_lens=txt_seq_lens)


def reference_gen_loop(sd, cfg: OracleConfig, latents, prompt_embeds, prompt_mask, timesteps,
                       step_size: float = 0.1, img_shapes=None, txt_seq_lens=None):
    """src/qwen_image_gen/model.py:117-135: same loop, no img_shapes / txt_seq_lens kwargs,
    ``latents - residual * 0.1``."""
    B = latents.shape[0]
    for t in timesteps:
        sample = transformer_forward(sd, cfg, latents.flatten(2).transpose(1, 2), prompt_embeds, prompt_mask,
                                     t.expand(B), img_shapes=img_shapes, txt_seq_lens=txt_seq_lens)
        residual = sample.transpose(1, 2).unflatten(2, latents.shape[2:])
        latents = latents - residual * step_size if step_size != 1.0 else latents - residual
    return latents


# --------------------------------------------------------------------------------------
# Synthetic inputs
# --------------------------------------------------------------------------------------
def synthetic_inputs(cfg: OracleConfig, height: int, width: int, s_txt: int, batch: int = 1,
                     dtype=torch.float32):
    Hl, Wl = latent_hw(height, width)
    g = torch.Generator().manual_seed(42)
    lat = torch.randn(batch, 1, 16, Hl, Wl, generator=g)
    g1 = torch.Generator().manual_seed(1)
    g2 = torch.Generator().manual_seed(2)
    cond = torch.randn(batch, s_txt, cfg.joint_attention_dim, generator=g1).to(dtype)
    uncond = torch.randn(batch, s_txt, cfg.joint_attention_dim, generator=g2).to(dtype)
    mask = torch.ones(batch, s_txt, dtype=torch.bool)
    packed = pack_latents(lat).to(dtype)
    img_shapes = [[(1, Hl // 2, Wl // 2)]] * batch
    return packed, cond, uncond, mask, img_shapes


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))
