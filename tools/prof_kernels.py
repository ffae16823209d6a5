"""Launches the hot kernels alone at the cfg-2 shapes, as the step issues them (batch 2 = cond + uncond stacked,
fp32 residual streams, the schedule the step uses: the persisted / measured (pair, tile width, rasterisation) of each
shape is registered first, ops.tune_gemm), for `ncu --set full`.
usage: prof_kernels.py [all|gemm|mlp_up|qkv|attn|elem] [batch]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops, rope  # noqa: E402

d, H, Dh, s_img, s_txt = 3072, 24, 128, 4096, 512
S = s_img + s_txt
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).bfloat16()
r32 = lambda *s, sc=1.0: torch.randn(*s, device=dev, generator=g) * sc
emp = lambda *s: torch.empty(*s, device=dev, dtype=torch.bfloat16)
which = sys.argv[1] if len(sys.argv) > 1 else "all"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 2
reps = 3


def launch(make):
    """make(i) -> the launch's problems; the shape's measured schedule is registered before the profiled launches."""
    ops.tune_gemm(make(0))
    torch.cuda.synchronize()
    for i in range(reps):
        ops.gemm(make(i))


xi, xt = r(B, s_img, d), r(B, s_txt, d)
h32, e32 = r32(B, s_img, d), r32(B, s_txt, d)              # fp32 residual streams
gate32 = r32(1, d, sc=0.1).expand(B, d)                    # fp32 gate row of the modulation table, shared by the batch
if which in ("all", "gemm", "mlp_up"):
    w_up = [r(2, 4 * d, d, sc=0.02) for _ in range(reps)]
    b_up = r(2, 4 * d)
    mi, mt = emp(B, s_img, 4 * d), emp(B, s_txt, 4 * d)
    launch(lambda i: [ops.gemm_problem(xi, w_up[i][0], b_up[0], mi, ops.EPI_BIAS_GELU_TANH),
                  ops.gemm_problem(xt, w_up[i][1], b_up[1], mt, ops.EPI_BIAS_GELU_TANH)])
if which in ("all", "gemm"):
    w_dn = [r(2, d, 4 * d, sc=0.02) for _ in range(reps)]
    launch(lambda i: [ops.gemm_problem(mi, w_dn[i][0], b_up[0, :d], h32, ops.EPI_GATED_RESIDUAL, h32, gate32),
                  ops.gemm_problem(mt, w_dn[i][1], b_up[1, :d], e32, ops.EPI_GATED_RESIDUAL, e32, gate32)])
    w_o = [r(2, d, d, sc=0.02) for _ in range(reps)]
    att = r(B, S, d)
    launch(lambda i: [ops.gemm_problem(att[:, s_txt:], w_o[i][0], b_up[0, :d], h32, ops.EPI_GATED_RESIDUAL, h32, gate32),
                  ops.gemm_problem(att[:, :s_txt], w_o[i][1], b_up[1, :d], e32, ops.EPI_GATED_RESIDUAL, e32, gate32)])
if which in ("all", "gemm", "qkv"):
    w_qkv = [r(2, 3 * d, d, sc=0.02) for _ in range(reps)]
    b_qkv = r(2, 3 * d)
    fr_ = rope.joint_cos_sin((16, 56, 56), 10000.0, [(1, 64, 64)], s_txt).to(dev)
    nw_ = r(4, Dh) + 1
    qkv_o = emp(B, S, 3 * d)
    kw = dict(epilogue=ops.EPI_QKV_NORM_ROPE, freqs=fr_, heads=H, head_dim=Dh)
    launch(lambda i: [ops.gemm_problem(xi, w_qkv[i][0], b_qkv[0], qkv_o[:, s_txt:], qk_norm=(nw_[0], nw_[1]), rope_pos0=s_txt, **kw),
                  ops.gemm_problem(xt, w_qkv[i][1], b_qkv[1], qkv_o[:, :s_txt], qk_norm=(nw_[2], nw_[3]), rope_pos0=0, **kw)])
if which in ("all", "attn"):
    qkv = r(B, S, 3 * d)
    q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
    o = emp(B, S, H, Dh)
    for i in range(reps):
        ops.attention(q, k, v, o)
if which in ("all", "elem"):
    mod32 = r32(1, 6 * d, sc=0.1)
    oi, ot = emp(B * s_img, d), emp(B * s_txt, d)
    lat = [r32(1, s_img, 64) for _ in range(3)]
    lat_b = emp(1, s_img, 64)
    for i in range(reps):
        ops.adaln_modulate_pair([(h32.view(B * s_img, d), mod32[:, :d].expand(B, d), mod32[:, d:2 * d].expand(B, d), oi, s_img),
                                 (e32.view(B * s_txt, d), mod32[:, :d].expand(B, d), mod32[:, d:2 * d].expand(B, d), ot, s_txt)])
        ops.cfg_euler_step_f32(lat[0], lat[1], lat[2], lat[2], lat_b, 4.0, dt=-0.02)
torch.cuda.synchronize()
print("ok")
