"""`torch.ops.qi.*` — the C ABI (include/qi_b200.h) registered with the PyTorch dispatcher.

SURVEY §8b: "Python side registers them as `torch.library` custom ops (so `inference_mode`, streams and CUDA graphs
work) and the `nn.Module` drop-in of boundary #1 calls those ops".  Each op is a thin CUDA-key implementation over
the ctypes binding in `ops.py` (one hand-written sm_100a kernel launch on the current stream) plus a fake-tensor
implementation, so the forward is traceable (`torch.compile(fullgraph=True)` sees opaque mutating ops) and usable
under `torch.inference_mode()` and CUDA-graph capture.  There is no CPU implementation: calling an op with CPU
tensors fails in the dispatcher ("no kernel for backend CPU"), the library itself raises without an sm_100 device.

All ops mutate caller-owned output tensors and return nothing — the C ABI's ownership rule (the caller owns every
buffer).  Lists hold the img-stream and txt-stream problem of one persistent launch.

Not registered: the peer-memory scatter variants (`qi_gemm_bf16` with `scatter`, `qi_attention_fwd_scatter`,
`qi_peer_barrier`) — their destinations are CUDA-IPC mapped raw pointers of other processes, which have no tensor
representation; the launcher calls them through `ops.py` directly.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import ops

_lib = torch.library.Library("qi", "DEF")

_SCHEMAS = {
    "gemm_bf16": "(Tensor[] a, Tensor[] w, Tensor?[] bias, Tensor(a!)[] out, int epilogue, Tensor?[] residual, "
                 "Tensor?[] gate, bool a_k_chunked=False) -> ()",
    "gemm_qkv_norm_rope": "(Tensor[] a, Tensor[] w, Tensor?[] bias, Tensor(a!)[] out, Tensor[] norm_wq, Tensor[] norm_wk, "
                          "Tensor freqs, int[] rope_pos0, int heads, int head_dim, float eps=1e-6) -> ()",
    "attention_fwd": "(Tensor q, Tensor k, Tensor v, Tensor(a!) out, float? softmax_scale=None) -> ()",
    "adaln_modulate": "(Tensor[] x, Tensor[] shift, Tensor[] scale, Tensor(a!)[] out, int[] rows_per_batch, "
                      "float eps=1e-6) -> ()",
    "rmsnorm_affine": "(Tensor x, Tensor weight, Tensor(a!) out, float eps=1e-6) -> ()",
    "qknorm_rope": "(Tensor(a!) qkv, int heads, int head_dim, int seq, int s_txt, Tensor wq_img, Tensor wk_img, "
                   "Tensor wq_txt, Tensor wk_txt, Tensor freqs, float eps=1e-6) -> ()",
    "cfg_euler_step": "(Tensor cond, Tensor? uncond, Tensor latents, Tensor(a!) out, float true_cfg_scale, "
                      "Tensor? dt_dev, float dt=0.0) -> ()",
    "cfg_euler_step_f32": "(Tensor cond, Tensor? uncond, Tensor latents, Tensor(a!) out, Tensor(b!)? out_bf16, "
                          "float true_cfg_scale, Tensor? dt_dev, float dt=0.0) -> ()",
    "small_linear": "(Tensor x, Tensor w, Tensor? bias, Tensor(a!) out, bool silu_in=False, bool silu_out=False) -> ()",
    "timestep_sinusoid": "(Tensor t, Tensor(a!) out) -> ()",
}
for _name, _schema in _SCHEMAS.items():
    _lib.define(_name + _schema)


def _gemm_bf16(a: List[torch.Tensor], w: List[torch.Tensor], bias: List[Optional[torch.Tensor]],
               out: List[torch.Tensor], epilogue: int, residual: List[Optional[torch.Tensor]],
               gate: List[Optional[torch.Tensor]], a_k_chunked: bool = False) -> None:
    n = len(a)
    if not (1 <= n <= 2) or any(len(x) != n for x in (w, bias, out, residual, gate)):
        raise ValueError("qi::gemm_bf16 takes 1 or 2 problems, every list of the same length")
    ops.gemm([ops.gemm_problem(a[i], w[i], bias[i], out[i], epilogue, residual[i], gate[i], a_k_chunked=a_k_chunked)
              for i in range(n)])


def _gemm_qkv_norm_rope(a, w, bias, out, norm_wq, norm_wk, freqs, rope_pos0, heads, head_dim, eps=1e-6) -> None:
    n = len(a)
    if not (1 <= n <= 2) or any(len(x) != n for x in (w, bias, out, norm_wq, norm_wk, rope_pos0)):
        raise ValueError("qi::gemm_qkv_norm_rope takes 1 or 2 problems, every list of the same length")
    ops.gemm([ops.gemm_problem(a[i], w[i], bias[i], out[i], ops.EPI_QKV_NORM_ROPE, qk_norm=(norm_wq[i], norm_wk[i]),
                               freqs=freqs, rope_pos0=int(rope_pos0[i]), heads=heads, head_dim=head_dim, eps=eps)
              for i in range(n)])


def _attention_fwd(q, k, v, out, softmax_scale=None) -> None:
    ops.attention(q, k, v, out, softmax_scale)


def _adaln_modulate(x, shift, scale, out, rows_per_batch, eps=1e-6) -> None:
    n = len(x)
    if not (1 <= n <= 2) or any(len(t) != n for t in (shift, scale, out, rows_per_batch)):
        raise ValueError("qi::adaln_modulate takes 1 or 2 row sets, every list of the same length")
    ops.adaln_modulate_pair([(x[i], shift[i], scale[i], out[i], int(rows_per_batch[i])) for i in range(n)], eps)


def _rmsnorm_affine(x, weight, out, eps=1e-6) -> None:
    ops.rmsnorm_affine(x, weight, out, eps)


def _qknorm_rope(qkv, heads, head_dim, seq, s_txt, wq_img, wk_img, wq_txt, wk_txt, freqs, eps=1e-6) -> None:
    ops.qknorm_rope(qkv, heads, head_dim, seq, s_txt, wq_img, wk_img, wq_txt, wk_txt, freqs, eps)


def _cfg_euler_step(cond, uncond, latents, out, true_cfg_scale, dt_dev, dt=0.0) -> None:
    ops.cfg_euler_step(cond, uncond, latents, out, true_cfg_scale, dt, dt_dev)


def _cfg_euler_step_f32(cond, uncond, latents, out, out_bf16, true_cfg_scale, dt_dev, dt=0.0) -> None:
    ops.cfg_euler_step_f32(cond, uncond, latents, out, out_bf16, true_cfg_scale, dt, dt_dev)


def _small_linear(x, w, bias, out, silu_in=False, silu_out=False) -> None:
    ops.small_linear(x, w, bias, out, silu_in, silu_out)


def _timestep_sinusoid(t, out) -> None:
    ops.timestep_sinusoid(t, out)


_IMPLS = {"gemm_bf16": _gemm_bf16, "gemm_qkv_norm_rope": _gemm_qkv_norm_rope, "attention_fwd": _attention_fwd,
          "adaln_modulate": _adaln_modulate, "rmsnorm_affine": _rmsnorm_affine, "qknorm_rope": _qknorm_rope,
          "cfg_euler_step": _cfg_euler_step, "cfg_euler_step_f32": _cfg_euler_step_f32, "small_linear": _small_linear, "timestep_sinusoid": _timestep_sinusoid}


def _fake(*args, **kwargs) -> None:
    """Every op writes into caller-provided outputs of caller-decided shape: nothing to infer."""
    return None


for _name, _fn in _IMPLS.items():
    _lib.impl(_name, _fn, "CUDA")
    torch.library.register_fake("qi::" + _name, _fake, lib=_lib)

OP_NAMES = tuple(_IMPLS)
