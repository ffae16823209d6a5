"""Drop-in for the transformer the reference calls every denoising step.

Boundary #1: ``transformer(hidden_states=, encoder_hidden_states=,
encoder_hidden_states_mask=, timestep=, img_shapes=, txt_seq_lens=, guidance=, return_dict=)``
-> object with ``.sample`` — call sites src/qwen_image/runner_adapter.py:128-140 and
src/qwen_image_gen/model.py:122-134 of the reference; the class it replaces is diffusers'
``QwenImageTransformer2DModel`` (operator list: src/doc/archive/QwenImage_Diffusers_Interface.md).

PyTorch is the host layer only: parameter storage, workspace allocation, stream.  Every FLOP of the
forward runs in libqi_b200.so (hand-written sm_100a kernels).  Per MMDiT block the launch sequence is

    K1     adaLN-LN-modulate (img + txt in one launch) -> xn
    K3+K2  QKV GEMM, img+txt in one persistent launch, QK-RMSNorm + MSRoPE applied per head vector in
           the epilogue -> joint qkv [B, S, 3d] (text rows first)
    K4     joint flash attention                    -> attn [B, S, d]
    K3     out-proj GEMM + gated residual (in place on h / e)
    K1 x2  adaLN-LN-modulate with mod2
    K3     MLP-up GEMM + GELU-tanh
    K3     MLP-down GEMM + gated residual (in place)

All 2L+1 modulation Linears share one stacked weight [L*12d + 2d, d]; one HBM-bound launch produces
the whole modulation vector of a forward, or — hoisted — one tcgen05 GEMM produces it for every
step of a schedule before the loop starts (temb depends only on t).
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from types import SimpleNamespace
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from . import ops, rope
from . import torch_ops  # noqa: F401  (registers torch.ops.qi.*)

Q = torch.ops.qi      # the C ABI as dispatcher-visible ops (torch_ops.py); `ops` is used directly only for the peer-scatter variants
from .config import MMDiTConfig


@dataclass
class Transformer2DModelOutput:
    sample: torch.Tensor


class _Workspace:
    def __init__(self, cfg: MMDiTConfig, B: int, s_img: int, s_txt: int, device, dtype,
                 residual_dtype=torch.float32):
        d = cfg.inner_dim
        S = s_img + s_txt
        e = lambda *shape: torch.empty(*shape, device=device, dtype=dtype)
        # residual streams: fp32 by default.  In bf16 the two in-place roundings per block accumulate like
        # sqrt(2 L) * 1.7e-3 = 1.9e-2 relative error over 60 blocks (measured, profiles/r02_depth_parity_bf16_stream.json)
        self.h = torch.empty(B, s_img, d, device=device, dtype=residual_dtype)
        self.e = torch.empty(B, s_txt, d, device=device, dtype=residual_dtype)
        self.xn_img = e(B, s_img, d)
        self.xn_txt = e(B, s_txt, d)
        self.txt_normed = e(B, s_txt, cfg.joint_attention_dim)
        self.qkv = e(B, S, 3 * d)
        self.attn = e(B, S, d)
        self.mlp_img = e(B, s_img, 4 * d)
        self.mlp_txt = e(B, s_txt, 4 * d)
        # timestep embedding and modulation follow the residual streams' precision (fp32: no rounding between
        # t and the shift / scale / gate vectors; they feed every block, so their error is coherent)
        md = residual_dtype
        self.sinus = torch.empty(B, 256, device=device, dtype=md)
        self.t1 = torch.empty(B, d, device=device, dtype=md)
        self.temb = torch.empty(B, d, device=device, dtype=md)
        self.t32 = torch.empty(B, device=device, dtype=torch.float32)
        self.sp = None

    def alloc_sp(self, cfg: MMDiTConfig, P: int, device, dtype):
        """NCCL Ulysses exchange buffers (B == 1): send/recv [P, S_loc, 3, H/P, Dh], attention output
        [P*S_loc, H/P, Dh], and the head-gathered result [P, S_loc, (H/P)*Dh]."""
        s_loc = self.qkv.shape[1]
        hs, dh = cfg.num_attention_heads // P, cfg.attention_head_dim
        e = lambda *shape: torch.empty(*shape, device=device, dtype=dtype)
        self.a2a_send = e(P, s_loc, 3, hs, dh)
        self.a2a_recv = e(P, s_loc, 3, hs, dh)
        self.o_full = e(P * s_loc, hs, dh)
        self.a2a_back = e(P, s_loc, hs * dh)
        self.sp = P


class B200QwenImageTransformer2DModel(nn.Module):
    def __init__(self, config: Optional[MMDiTConfig] = None, device="cuda", dtype=torch.bfloat16):
        super().__init__()
        if dtype != torch.bfloat16:
            raise ValueError("the B200 hot path computes in bf16 (fp32 accumulate) only")
        self.cfg = config or MMDiTConfig()
        cfg = self.cfg
        d, L = cfg.inner_dim, cfg.num_layers
        self.config = SimpleNamespace(in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                                      guidance_embeds=cfg.guidance_embeds, num_layers=L,
                                      attention_head_dim=cfg.attention_head_dim,
                                      num_attention_heads=cfg.num_attention_heads,
                                      joint_attention_dim=cfg.joint_attention_dim,
                                      patch_size=cfg.patch_size, axes_dims_rope=cfg.axes_dims_rope)
        self._attention_backend = "qi_tcgen05_flash"
        # QK-RMSNorm + MSRoPE inside the QKV GEMM epilogue (default) or as the separate K2 launch
        self.fuse_qknorm_rope = os.environ.get("QI_FUSE_QKNORM", "1") != "0"
        # dtype of the residual streams h / e between blocks (QI_RESIDUAL=bf16 keeps upstream's bf16 storage: A/B aid)
        self.residual_dtype = torch.bfloat16 if os.environ.get("QI_RESIDUAL", "fp32") == "bf16" else torch.float32

        def P(*shape):
            return nn.Parameter(torch.empty(*shape, device=device, dtype=dtype), requires_grad=False)

        self.img_in_w, self.img_in_b = P(d, cfg.in_channels), P(d)
        self.txt_norm_w = P(cfg.joint_attention_dim)
        self.txt_in_w, self.txt_in_b = P(d, cfg.joint_attention_dim), P(d)
        self.time_w1, self.time_b1 = P(d, 256), P(d)
        self.time_w2, self.time_b2 = P(d, d), P(d)
        # stacked modulation: rows [(l*2+s)*6d, +6d) for block l stream s (0 img, 1 txt); last 2d = norm_out
        self.n_mod = L * 12 * d + 2 * d
        self.mod_w, self.mod_b = P(self.n_mod, d), P(self.n_mod)
        self.qkv_w = P(L, 2, 3 * d, d)      # [block, stream, (q|k|v) out, in]
        self.qkv_b = P(L, 2, 3 * d)
        self.qk_norm_w = P(L, 4, cfg.attention_head_dim)   # norm_q, norm_k, norm_added_q, norm_added_k
        self.out_w, self.out_b = P(L, 2, d, d), P(L, 2, d)
        self.mlp_up_w, self.mlp_up_b = P(L, 2, 4 * d, d), P(L, 2, 4 * d)
        self.mlp_down_w, self.mlp_down_b = P(L, 2, d, 4 * d), P(L, 2, d)
        self.proj_out_w, self.proj_out_b = P(cfg.proj_out_dim, d), P(cfg.proj_out_dim)
        self._ws: Dict[Tuple[int, int, int], _Workspace] = {}
        self._freqs: Dict[tuple, torch.Tensor] = {}
        self._mod_table: Optional[torch.Tensor] = None      # [steps, n_mod] hoisted modulation
        # bumped by every weight (re)load: whatever was derived from the weights (hoisted modulation tables, CUDA
        # graphs captured over them) is keyed on state_signature() by its owner (pipeline sessions, launcher setup)
        self.weights_generation = 0
        self._mod_steps: Optional[torch.Tensor] = None
        self._sp = None              # UlyssesContext when sequence-parallel
        self._sp_freqs: Optional[torch.Tensor] = None

    # ------------------------------------------------------------------ sequence parallelism (Ulysses)
    def set_sequence_parallel(self, ctx) -> None:
        """ctx: launcher.UlyssesContext (group, size, rank).  Heads must divide by the degree."""
        if ctx is not None and self.cfg.num_attention_heads % ctx.size != 0:
            raise ValueError("num_attention_heads must be divisible by the sequence-parallel degree")
        self._sp = ctx
        self._ws.clear()

    def set_sequence_shard(self, img_shape, s_img_full: int, s_txt_full: int, img_range, txt_range,
                           ref_shape=None, ref_range=None) -> None:
        """This rank owns image tokens [i0, i1) and text tokens [t0, t1) of the full sequence — and, in edit
        mode, tokens [j0, j1) of the reference image (second entry of img_shapes, RoPE frame index 1); its RoPE
        rows are the matching rows of the full joint table (text rows first, then image 0, then image 1)."""
        shapes = [tuple(int(v) for v in img_shape)]
        if ref_shape is not None:
            shapes.append(tuple(int(v) for v in ref_shape))
        full = rope.joint_cos_sin(self.cfg.axes_dims_rope, self.cfg.rope_theta, tuple(shapes), s_txt_full)
        (i0, i1), (t0, t1) = img_range, txt_range
        parts = [full[t0:t1], full[s_txt_full + i0:s_txt_full + i1]]
        if ref_shape is not None:
            j0, j1 = ref_range
            r0 = s_txt_full + s_img_full
            parts.append(full[r0 + j0:r0 + j1])
        self._sp_freqs = torch.cat(parts, dim=0).contiguous().to(self.device)

    # ------------------------------------------------------------------ host-layer conveniences
    @property
    def dtype(self):
        return self.img_in_w.dtype

    @property
    def device(self):
        return self.img_in_w.device

    def set_attention_backend(self, name: str) -> None:
        """Reference selects '_flash_3_hub' (test_performance.py:324); here there is exactly one
        attention kernel (qi_attention_fwd), so the name is recorded and otherwise ignored."""
        self._attention_backend = str(name)

    @classmethod
    def from_pretrained(cls, model_id, subfolder: Optional[str] = None, torch_dtype=torch.bfloat16,
                        device_map=None, config: Optional[MMDiTConfig] = None, device="cuda"):
        """Host-layer weight loading (reference: model.py:52-57).  Accepts a local directory holding
        diffusers-named ``*.safetensors`` shards; there is no network access to a hub."""
        import glob
        import os
        path = os.path.join(model_id, subfolder) if subfolder else model_id
        files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
        if not files:
            raise RuntimeError(f"Failed to load QwenImage components: no safetensors under {path!r}")
        try:
            from safetensors.torch import load_file
        except ImportError as e:  # pragma: no cover
            raise ImportError("safetensors is required to load checkpoints") from e
        sd = {}
        for f in files:
            sd.update(load_file(f))
        m = cls(config, device=device, dtype=torch_dtype)
        m.load_diffusers_state_dict(sd)
        return m

    @torch.no_grad()
    def load_diffusers_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        """Pack a diffusers-named state dict into the fused layout."""
        cfg, d = self.cfg, self.cfg.inner_dim
        cp = lambda dst, src: dst.copy_(src.to(dst.dtype))
        cp(self.img_in_w, sd["img_in.weight"]); cp(self.img_in_b, sd["img_in.bias"])
        cp(self.txt_norm_w, sd["txt_norm.weight"])
        cp(self.txt_in_w, sd["txt_in.weight"]); cp(self.txt_in_b, sd["txt_in.bias"])
        t = "time_text_embed.timestep_embedder."
        cp(self.time_w1, sd[t + "linear_1.weight"]); cp(self.time_b1, sd[t + "linear_1.bias"])
        cp(self.time_w2, sd[t + "linear_2.weight"]); cp(self.time_b2, sd[t + "linear_2.bias"])
        for l in range(cfg.num_layers):
            p = f"transformer_blocks.{l}."
            for s, name in enumerate(("img_mod.1", "txt_mod.1")):
                r0 = (l * 2 + s) * 6 * d
                cp(self.mod_w[r0:r0 + 6 * d], sd[p + name + ".weight"])
                cp(self.mod_b[r0:r0 + 6 * d], sd[p + name + ".bias"])
            for s, names in enumerate((("to_q", "to_k", "to_v"), ("add_q_proj", "add_k_proj", "add_v_proj"))):
                for j, n in enumerate(names):
                    cp(self.qkv_w[l, s, j * d:(j + 1) * d], sd[p + "attn." + n + ".weight"])
                    cp(self.qkv_b[l, s, j * d:(j + 1) * d], sd[p + "attn." + n + ".bias"])
            for j, n in enumerate(("norm_q", "norm_k", "norm_added_q", "norm_added_k")):
                cp(self.qk_norm_w[l, j], sd[p + "attn." + n + ".weight"])
            for s, n in enumerate(("attn.to_out.0", "attn.to_add_out")):
                cp(self.out_w[l, s], sd[p + n + ".weight"]); cp(self.out_b[l, s], sd[p + n + ".bias"])
            for s, n in enumerate(("img_mlp", "txt_mlp")):
                cp(self.mlp_up_w[l, s], sd[p + n + ".net.0.proj.weight"])
                cp(self.mlp_up_b[l, s], sd[p + n + ".net.0.proj.bias"])
                cp(self.mlp_down_w[l, s], sd[p + n + ".net.2.weight"])
                cp(self.mlp_down_b[l, s], sd[p + n + ".net.2.bias"])
        r0 = cfg.num_layers * 12 * d
        cp(self.mod_w[r0:], sd["norm_out.linear.weight"]); cp(self.mod_b[r0:], sd["norm_out.linear.bias"])
        cp(self.proj_out_w, sd["proj_out.weight"]); cp(self.proj_out_b, sd["proj_out.bias"])
        self._mod_table = None
        self.weights_generation += 1

    def diffusers_param(self, name: str) -> torch.Tensor:
        """Inverse of `load_diffusers_state_dict` for ONE diffusers-named tensor: a view
        into the fused parameter that holds it.  Lets a checker (or an exporter) walk the 20B model block by
        block without materialising a second copy of the weights."""
        d = self.cfg.inner_dim
        top = {"img_in.weight": self.img_in_w, "img_in.bias": self.img_in_b, "txt_norm.weight": self.txt_norm_w,
               "txt_in.weight": self.txt_in_w, "txt_in.bias": self.txt_in_b,
               "time_text_embed.timestep_embedder.linear_1.weight": self.time_w1,
               "time_text_embed.timestep_embedder.linear_1.bias": self.time_b1,
               "time_text_embed.timestep_embedder.linear_2.weight": self.time_w2,
               "time_text_embed.timestep_embedder.linear_2.bias": self.time_b2,
               "proj_out.weight": self.proj_out_w, "proj_out.bias": self.proj_out_b}
        if name in top:
            return top[name]
        r0 = self.cfg.num_layers * 12 * d
        if name == "norm_out.linear.weight":
            return self.mod_w[r0:]
        if name == "norm_out.linear.bias":
            return self.mod_b[r0:]
        parts = name.split(".")
        if parts[0] != "transformer_blocks":
            raise KeyError(name)
        l, rest, kind = int(parts[1]), ".".join(parts[2:-1]), parts[-1]
        wb = lambda w, b: w if kind == "weight" else b
        if rest in ("img_mod.1", "txt_mod.1"):
            m0 = (l * 2 + (rest == "txt_mod.1")) * 6 * d
            return wb(self.mod_w, self.mod_b)[m0:m0 + 6 * d]
        qkv = {"attn.to_q": (0, 0), "attn.to_k": (0, 1), "attn.to_v": (0, 2),
               "attn.add_q_proj": (1, 0), "attn.add_k_proj": (1, 1), "attn.add_v_proj": (1, 2)}
        if rest in qkv:
            s, j = qkv[rest]
            return wb(self.qkv_w, self.qkv_b)[l, s, j * d:(j + 1) * d]
        norms = {"attn.norm_q": 0, "attn.norm_k": 1, "attn.norm_added_q": 2, "attn.norm_added_k": 3}
        if rest in norms:
            return self.qk_norm_w[l, norms[rest]]
        outs = {"attn.to_out.0": 0, "attn.to_add_out": 1}
        if rest in outs:
            return wb(self.out_w, self.out_b)[l, outs[rest]]
        mlps = {"img_mlp.net.0.proj": (0, 0), "txt_mlp.net.0.proj": (1, 0), "img_mlp.net.2": (0, 1),
                "txt_mlp.net.2": (1, 1)}
        if rest in mlps:
            s, dn = mlps[rest]
            return (wb(self.mlp_down_w, self.mlp_down_b) if dn else wb(self.mlp_up_w, self.mlp_up_b))[l, s]
        raise KeyError(name)

    @torch.no_grad()
    def init_random_(self, seed: int = 0, weight_std: float = 0.02, bias_std: float = 0.0,
                     norm_weight_std: float = 0.0) -> "B200QwenImageTransformer2DModel":
        """Random-init on the device for benchmarking at 20B scale (Linear.weight
        ~ N(0, 0.02^2), bias 0, norm weights 1).  Uses the device generator, not the CPU one.
        bias_std / norm_weight_std > 0 (parity tests) exercise the bias and affine-norm paths."""
        g = torch.Generator(device=self.device).manual_seed(seed)
        for name, p in self.named_parameters():
            if name in ("txt_norm_w", "qk_norm_w"):
                p.fill_(1.0)
                if norm_weight_std > 0:
                    p.add_((torch.randn(p.shape, device=self.device, generator=g) * norm_weight_std).to(p.dtype))
            elif name.endswith("_b"):
                p.zero_()
                if bias_std > 0:
                    p.copy_((torch.randn(p.shape, device=self.device, generator=g) * bias_std).to(p.dtype))
            else:
                flat = p.view(-1)
                step = 1 << 28
                for i in range(0, flat.numel(), step):
                    n = min(step, flat.numel() - i)
                    flat[i:i + n] = (torch.randn(n, device=self.device, generator=g) * weight_std).to(p.dtype)
        self._mod_table = None
        self.weights_generation += 1
        return self

    def state_signature(self) -> tuple:
        """Everything a cached denoising setup depends on besides shapes and schedule: the weights' generation and
        the two switches that change which kernels / buffer dtypes a captured step uses.  A cache whose signature
        differs from this must be rebuilt (in-place edits of the parameters that bypass load_diffusers_state_dict /
        init_random_ need `weights_generation += 1` by hand)."""
        return (self.weights_generation, bool(self.fuse_qknorm_rope), str(self.residual_dtype))

    # ------------------------------------------------------------------ GEMM schedule autotuning
    @torch.no_grad()
    def tune_schedules(self, B: int, s_img: int, s_txt: int) -> dict:
        """Time every GEMM schedule (single-CTA / CTA-pair x tile width) for the four block GEMMs at this
        forward's shapes and register the fastest with the library (ops.tune_gemm).  Uses the forward's
        own workspace as scratch (every buffer is rewritten by the next forward); call before the step is
        captured into a CUDA graph.  Tile quantisation over 148 SMs makes the best schedule shape
        dependent (sequence-parallel shards especially), which a static heuristic only approximates."""
        if os.environ.get("QI_GEMM_TUNE", "1") == "0":
            return {}
        cfg, dev, dt = self.cfg, self.device, self.dtype
        d, H, Dh = cfg.inner_dim, cfg.num_attention_heads, cfg.attention_head_dim
        key = (B, s_img, s_txt)
        ws = self._ws.get(key)
        if ws is None:
            ws = self._ws[key] = _Workspace(cfg, B, s_img, s_txt, dev, dt, self.residual_dtype)
        for t in (ws.h, ws.e, ws.xn_img, ws.xn_txt, ws.attn, ws.mlp_img, ws.mlp_txt):
            t.zero_()
        S = s_img + s_txt
        freqs = torch.zeros(S, Dh // 2, 2, device=dev, dtype=torch.float32)
        gate = torch.zeros(B, d, device=dev, dtype=self.residual_dtype)   # same dtype as the modulation table
        nw = self.qk_norm_w[0]
        G = ops.EPI_GATED_RESIDUAL
        qkv_img, qkv_txt = ws.qkv[:, s_txt:], ws.qkv[:, :s_txt]
        att_img, att_txt = ws.attn[:, s_txt:], ws.attn[:, :s_txt]
        chosen = {}
        if self.fuse_qknorm_rope:
            qk = dict(epilogue=ops.EPI_QKV_NORM_ROPE, freqs=freqs, heads=H, head_dim=Dh)
            probs = [ops.gemm_problem(ws.xn_img, self.qkv_w[0, 0], self.qkv_b[0, 0], qkv_img, qk_norm=(nw[0], nw[1]),
                                      rope_pos0=s_txt, **qk),
                     ops.gemm_problem(ws.xn_txt, self.qkv_w[0, 1], self.qkv_b[0, 1], qkv_txt, qk_norm=(nw[2], nw[3]),
                                      rope_pos0=0, **qk)]
        else:
            probs = [ops.gemm_problem(ws.xn_img, self.qkv_w[0, 0], self.qkv_b[0, 0], qkv_img),
                     ops.gemm_problem(ws.xn_txt, self.qkv_w[0, 1], self.qkv_b[0, 1], qkv_txt)]
        chosen["qkv"] = ops.tune_gemm(probs)
        chosen["out"] = ops.tune_gemm([
            ops.gemm_problem(att_img, self.out_w[0, 0], self.out_b[0, 0], ws.h, G, ws.h, gate),
            ops.gemm_problem(att_txt, self.out_w[0, 1], self.out_b[0, 1], ws.e, G, ws.e, gate)])
        chosen["mlp_up"] = ops.tune_gemm([
            ops.gemm_problem(ws.xn_img, self.mlp_up_w[0, 0], self.mlp_up_b[0, 0], ws.mlp_img, ops.EPI_BIAS_GELU_TANH),
            ops.gemm_problem(ws.xn_txt, self.mlp_up_w[0, 1], self.mlp_up_b[0, 1], ws.mlp_txt, ops.EPI_BIAS_GELU_TANH)])
        chosen["mlp_down"] = ops.tune_gemm([
            ops.gemm_problem(ws.mlp_img, self.mlp_down_w[0, 0], self.mlp_down_b[0, 0], ws.h, G, ws.h, gate),
            ops.gemm_problem(ws.mlp_txt, self.mlp_down_w[0, 1], self.mlp_down_b[0, 1], ws.e, G, ws.e, gate)])
        torch.cuda.synchronize(dev)
        chosen["split_launches"] = len(ops._split_keys)     # two-problem dims issued as one launch per problem
        return chosen

    # ------------------------------------------------------------------ modulation
    def _time_embed(self, t32: torch.Tensor, ws_sinus, ws_t1, ws_temb) -> torch.Tensor:
        Q.timestep_sinusoid(t32, ws_sinus)
        Q.small_linear(ws_sinus, self.time_w1, self.time_b1, ws_t1, False, True)
        Q.small_linear(ws_t1, self.time_w2, self.time_b2, ws_temb)
        return ws_temb

    @torch.no_grad()
    def prepare_schedule(self, timesteps_over_1000: torch.Tensor) -> torch.Tensor:
        """Hoist every modulation Linear out of the denoising loop: temb depends only on t, so the
        [steps, L*12d+2d] modulation table is one tcgen05 GEMM over all steps.  `timesteps_over_1000`
        holds the values the pipeline would pass as `timestep` (already rounded to bf16)."""
        t = timesteps_over_1000.to(self.device, torch.bfloat16).to(torch.float32).contiguous()
        n, d = t.numel(), self.cfg.inner_dim
        md = self.residual_dtype
        e = lambda *s: torch.empty(*s, device=self.device, dtype=md)
        sinus, t1, temb = e(n, 256), e(n, d), e(n, d)
        self._time_embed(t, sinus, t1, temb)
        table = e(n, self.n_mod)
        if md == torch.float32:
            # fp32 end to end (weights bf16): HBM-bound on the 6.8 GB of modulation weights, once per schedule
            Q.small_linear(temb, self.mod_w, self.mod_b, table, True, False)
        else:
            silu = torch.nn.functional.silu(temb)   # host-layer glue on [steps, d]; 0.3 MB, once per image
            Q.gemm_bf16([silu], [self.mod_w], [self.mod_b], [table], ops.EPI_BIAS, [None], [None])
        self._mod_table, self._mod_steps = table, t
        return table

    def clear_schedule(self) -> None:
        self._mod_table = self._mod_steps = None

    def _modulation(self, ws: _Workspace, timestep: torch.Tensor, B: int, step_index: Optional[int]):
        """Returns (mod [rows, n_mod] view, batch_stride) — batch_stride 0 broadcasts one row."""
        if step_index is not None and self._mod_table is not None:
            return self._mod_table[step_index:step_index + 1], 0
        ws.t32.copy_(timestep.to(self.device).to(self.dtype).to(torch.float32))   # upstream rounds t to bf16
        temb = self._time_embed(ws.t32, ws.sinus, ws.t1, ws.temb)
        key = ("mod", B)
        if key not in self._freqs:
            self._freqs[key] = torch.empty(B, self.n_mod, device=self.device, dtype=self.residual_dtype)
        mod = self._freqs[key]
        Q.small_linear(temb, self.mod_w, self.mod_b, mod, True, False)
        return mod, self.n_mod

    # ------------------------------------------------------------------ forward
    @torch.no_grad()
    def forward(self, hidden_states: torch.Tensor, encoder_hidden_states: torch.Tensor = None,
                encoder_hidden_states_mask: torch.Tensor = None, timestep: torch.Tensor = None,
                img_shapes=None, txt_seq_lens=None, guidance: torch.Tensor = None,
                attention_kwargs=None, controlnet_block_samples=None, return_dict: bool = True,
                _step_index: Optional[int] = None, _mod_override: Optional[torch.Tensor] = None,
                _out_rows: Optional[int] = None, _block_hook=None, _out_fp32: bool = False):
        """Private kwargs (host layer of this package only): `_step_index` / `_mod_override` select a row of
        the hoisted modulation table; `_out_rows` = n computes `proj_out` for the first n image tokens of every
        sample only (edit mode keeps `noise_pred[:, :S_noise]`, so the reference-image rows need no projection);
        `_out_fp32` returns `.sample` in fp32 (the pipeline's fp32 CFG / Euler step reads it unrounded);
        `_block_hook(l, h, e)` is called after block l with the residual streams (parity tests log the error
        growth over depth with it)."""
        cfg = self.cfg
        if hidden_states.dim() != 3 or hidden_states.shape[-1] != cfg.in_channels:
            raise ValueError(f"hidden_states must be [B, S_img, {cfg.in_channels}]")
        if encoder_hidden_states is None or encoder_hidden_states.dim() != 3 \
                or encoder_hidden_states.shape[-1] != cfg.joint_attention_dim:
            raise ValueError(f"encoder_hidden_states must be [B, S_txt, {cfg.joint_attention_dim}]")
        if controlnet_block_samples is not None:
            raise NotImplementedError("controlnet residuals are outside the hot path")
        if timestep is None and _step_index is None and _mod_override is None:
            raise ValueError("timestep is required")
        B, s_img, _ = hidden_states.shape
        s_txt = encoder_hidden_states.shape[1]
        if encoder_hidden_states.shape[0] != B:
            raise ValueError("Batch size mismatch between latents and prompt embeddings")
        d, H, Dh, L = cfg.inner_dim, cfg.num_attention_heads, cfg.attention_head_dim, cfg.num_layers
        S = s_txt + s_img
        if img_shapes is None:
            side = int(round(math.sqrt(s_img)))
            if side * side != s_img:
                raise ValueError("img_shapes is required for non-square token grids")
            shapes = ((1, side, side),)
        else:
            shapes = rope.normalize_img_shapes(img_shapes)
        sp = self._sp
        if sp is None and sum(f * h * w for f, h, w in shapes) != s_img:
            raise ValueError("img_shapes does not match the number of image tokens")
        peer = sp.peer if sp is not None else None
        if sp is not None:
            if peer is None and B != 1:
                raise NotImplementedError("the NCCL Ulysses exchange is implemented for batch 1; "
                                          "use the peer-memory exchange for batched runs")
            if peer is not None and (peer.batch != B or peer.s_loc != s_img + s_txt):
                raise ValueError("peer exchange buffers were sized for a different batch / shard")
            if self._sp_freqs is None or self._sp_freqs.shape[0] != s_img + s_txt:
                raise ValueError("call set_sequence_shard() before a sequence-parallel forward")
        max_txt = int(max(txt_seq_lens)) if txt_seq_lens is not None else s_txt
        if max_txt != s_txt:
            raise ValueError("max(txt_seq_lens) must equal the text sequence length")
        dev, dt = self.device, self.dtype
        x_in = hidden_states.to(dev, dt).contiguous()
        enc = encoder_hidden_states.to(dev, dt).contiguous()

        key = (B, s_img, s_txt)
        ws = self._ws.get(key)
        if ws is None:
            ws = self._ws[key] = _Workspace(cfg, B, s_img, s_txt, dev, dt, self.residual_dtype)
        if sp is not None:
            freqs = self._sp_freqs
            if peer is None and ws.sp != sp.size:
                ws.alloc_sp(cfg, sp.size, dev, dt)
        else:
            fkey = (shapes, s_txt)
            freqs = self._freqs.get(fkey)
            if freqs is None:
                freqs = self._freqs[fkey] = rope.joint_cos_sin(cfg.axes_dims_rope, cfg.rope_theta, shapes, s_txt).to(dev)

        if _mod_override is not None:
            mod, mbs = _mod_override, 0
        else:
            mod, mbs = self._modulation(ws, timestep, B, _step_index)

        def mod_slices(row0):
            # (shift, scale, gate) views [Bm, d] of one adaLN triple starting at column row0
            return mod[:, row0:row0 + d], mod[:, row0 + d:row0 + 2 * d], mod[:, row0 + 2 * d:row0 + 3 * d]

        def bview(t):  # [1 or B, d] view whose batch stride is 0 when the row is shared
            return t if mbs != 0 else t.expand(B, d)

        # input projections
        Q.rmsnorm_affine(enc.view(B * s_txt, -1), self.txt_norm_w, ws.txt_normed.view(B * s_txt, -1))
        Q.gemm_bf16([x_in, ws.txt_normed], [self.img_in_w, self.txt_in_w], [self.img_in_b, self.txt_in_b],
                    [ws.h, ws.e], ops.EPI_BIAS, [None, None], [None, None])
        h, e = ws.h, ws.e
        h2, e2 = h.view(B * s_img, d), e.view(B * s_txt, d)
        xi2, xt2 = ws.xn_img.view(B * s_img, d), ws.xn_txt.view(B * s_txt, d)
        qkv_img, qkv_txt = ws.qkv[:, s_txt:], ws.qkv[:, :s_txt]
        att_img, att_txt = ws.attn[:, s_txt:], ws.attn[:, :s_txt]
        q = ws.qkv[:, :, 0:d].unflatten(-1, (H, Dh))
        k = ws.qkv[:, :, d:2 * d].unflatten(-1, (H, Dh))
        v = ws.qkv[:, :, 2 * d:3 * d].unflatten(-1, (H, Dh))
        attn4 = ws.attn.unflatten(-1, (H, Dh))
        G = ops.EPI_GATED_RESIDUAL
        N2 = [None, None]

        def gated(a_i, a_t, w, bias, g_i, g_t, chunked=False, wait=None):
            # out-projection / MLP-down of both streams: h, e updated in place by the gated-residual epilogue
            if wait is not None:    # A was stored by peers: the flags ride in the problem struct (ctypes call)
                ops.gemm([ops.gemm_problem(a_i, w[0], bias[0], h, G, h, bview(g_i), a_k_chunked=chunked, wait=wait),
                          ops.gemm_problem(a_t, w[1], bias[1], e, G, e, bview(g_t), a_k_chunked=chunked)])
                return
            Q.gemm_bf16([a_i, a_t], [w[0], w[1]], [bias[0], bias[1]], [h, e], G, [h, e],
                        [bview(g_i), bview(g_t)], chunked)

        for l in range(L):
            bi, bt = (l * 2) * 6 * d, (l * 2 + 1) * 6 * d
            sh_i1, sc_i1, g_i1 = mod_slices(bi)
            sh_i2, sc_i2, g_i2 = mod_slices(bi + 3 * d)
            sh_t1, sc_t1, g_t1 = mod_slices(bt)
            sh_t2, sc_t2, g_t2 = mod_slices(bt + 3 * d)
            Q.adaln_modulate([h2, e2], [bview(sh_i1), bview(sh_t1)], [bview(sc_i1), bview(sc_t1)], [xi2, xt2],
                             [s_img, s_txt])
            # QKV projection with QK-RMSNorm + MSRoPE applied per head vector in the GEMM epilogue (K2 fused
            # into K3); text rows first in the joint sequence, so the img stream's row 0 sits at table row s_txt
            nw = self.qk_norm_w[l]
            sc = None
            if peer is not None:
                sc = peer.dst_qkv if (peer.fold and self.fuse_qknorm_rope) else peer.dst_qkv_plain
            w_qkv, b_qkv = self.qkv_w[l], self.qkv_b[l]
            if self.fuse_qknorm_rope and sc is not None:
                # peer scatter: the head vectors leave through CUDA-IPC pointers (no tensor form) -> ctypes call
                qk = dict(epilogue=ops.EPI_QKV_NORM_ROPE, freqs=freqs, heads=H, head_dim=Dh, scatter=sc)
                ops.gemm([ops.gemm_problem(ws.xn_img, w_qkv[0], b_qkv[0], None, qk_norm=(nw[0], nw[1]),
                                           rope_pos0=s_txt, **qk),
                          ops.gemm_problem(ws.xn_txt, w_qkv[1], b_qkv[1], None, qk_norm=(nw[2], nw[3]),
                                           rope_pos0=0, **qk)])
            elif self.fuse_qknorm_rope:
                Q.gemm_qkv_norm_rope([ws.xn_img, ws.xn_txt], [w_qkv[0], w_qkv[1]], [b_qkv[0], b_qkv[1]],
                                     [qkv_img, qkv_txt], [nw[0], nw[2]], [nw[1], nw[3]], freqs, [s_txt, 0], H, Dh)
            else:   # separate K2 launch (QI_FUSE_QKNORM=0): same arithmetic, one more pass over qkv
                Q.gemm_bf16([ws.xn_img, ws.xn_txt], [w_qkv[0], w_qkv[1]], [b_qkv[0], b_qkv[1]], [qkv_img, qkv_txt],
                            ops.EPI_BIAS, N2, N2)
                if sc is not None:
                    ops.qknorm_rope_scatter(ws.qkv.view(B * S, 3 * d), H, Dh, S, s_txt, nw[0], nw[1], nw[2], nw[3],
                                            freqs, sc)
                else:
                    Q.qknorm_rope(ws.qkv.view(B * S, 3 * d), H, Dh, S, s_txt, nw[0], nw[1], nw[2], nw[3], freqs)
            if sp is None:
                Q.attention_fwd(q, k, v, attn4)
                gated(att_img, att_txt, self.out_w[l], self.out_b[l], g_i1, g_t1)
            elif peer is not None:
                # Fused exchange over peer memory: the QKV GEMM's epilogue has stored each head vector into
                # the head owner's HBM, attention stores each output row into the token owner's HBM.  The
                # "all peers' stores have landed" flags are folded into the kernels (producer's last CTA
                # signals, consumer's TMA thread polls: peer.sig_qkv / peer.sig_o inside the descriptors);
                # without the fused QKV epilogue, or with QI_PEER_FOLD=0, two flag-barrier kernels instead.
                folded = peer.fold and self.fuse_qknorm_rope
                dst_o = peer.dst_o if folded else peer.dst_o_plain
                if not folded:
                    peer.barrier()
                full = peer.recv                       # head-major [B, 3, H/P, S_total, Dh]
                ops.attention_scatter(full[:, 0].transpose(1, 2), full[:, 1].transpose(1, 2),
                                      full[:, 2].transpose(1, 2), dst_o, peer.o_ld, peer.o_batch_stride,
                                      o_head_stride=peer.o_head_stride)
                if not folded:
                    peer.barrier()
                back = peer.back[:, :, :S]             # [P*H/P, B, S_loc, Dh]: K-chunked A operand (chunk = head)
                gated(back[:, :, s_txt:], back[:, :, :s_txt], self.out_w[l], self.out_b[l], g_i1, g_t1, True,
                      wait=peer.sig_o if folded else None)
            else:
                # NCCL head<->sequence all-to-all: [S_loc, H] -> [S, H/P]; attention; back as K-chunked A
                full = sp.scatter_heads(ws.qkv.view(S, 3 * d), H, Dh, ws.a2a_send, ws.a2a_recv)
                Q.attention_fwd(full[:, 0].unsqueeze(0), full[:, 1].unsqueeze(0), full[:, 2].unsqueeze(0),
                                ws.o_full.unsqueeze(0))
                back = sp.gather_heads(ws.o_full, ws.a2a_back)
                gated(back[:, s_txt:], back[:, :s_txt], self.out_w[l], self.out_b[l], g_i1, g_t1, True)
            Q.adaln_modulate([h2, e2], [bview(sh_i2), bview(sh_t2)], [bview(sc_i2), bview(sc_t2)], [xi2, xt2],
                             [s_img, s_txt])
            Q.gemm_bf16([ws.xn_img, ws.xn_txt], [self.mlp_up_w[l, 0], self.mlp_up_w[l, 1]],
                        [self.mlp_up_b[l, 0], self.mlp_up_b[l, 1]], [ws.mlp_img, ws.mlp_txt],
                        ops.EPI_BIAS_GELU_TANH, N2, N2)
            gated(ws.mlp_img, ws.mlp_txt, self.mlp_down_w[l], self.mlp_down_b[l], g_i2, g_t2)
            if _block_hook is not None:
                _block_hook(l, h, e)
        # norm_out (AdaLayerNormContinuous: chunk order is (scale, shift)) + proj_out
        r0 = L * 12 * d
        Q.adaln_modulate([h2], [bview(mod[:, r0 + d:r0 + 2 * d])], [bview(mod[:, r0:r0 + d])], [xi2], [s_img])
        n_out = s_img if _out_rows is None else int(_out_rows)
        if not 0 < n_out <= s_img:
            raise ValueError("_out_rows must be in (0, S_img]")
        out = torch.empty(B, n_out, cfg.proj_out_dim, device=dev, dtype=torch.float32 if _out_fp32 else dt)
        Q.gemm_bf16([ws.xn_img[:, :n_out]], [self.proj_out_w], [self.proj_out_b], [out], ops.EPI_BIAS, [None], [None])
        if not return_dict:
            return (out,)
        return Transformer2DModelOutput(sample=out)
