"""Launches the hot kernels alone at the cfg-2 shapes (for `ncu --set full`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops  # noqa: E402

d, H, Dh, s_img, s_txt = 3072, 24, 128, 4096, 512
S = s_img + s_txt
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).bfloat16()
which = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = 3
xi, xt = r(1, s_img, d), r(1, s_txt, d)
if which in ("all", "gemm"):
    w_up = [r(2, 4 * d, d, sc=0.02) for _ in range(reps)]
    b_up = r(2, 4 * d)
    mi, mt = torch.empty(1, s_img, 4 * d, device=dev, dtype=torch.bfloat16), torch.empty(1, s_txt, 4 * d, device=dev, dtype=torch.bfloat16)
    for i in range(reps):
        ops.gemm([ops.gemm_problem(xi, w_up[i][0], b_up[0], mi, ops.EPI_BIAS_GELU_TANH),
                  ops.gemm_problem(xt, w_up[i][1], b_up[1], mt, ops.EPI_BIAS_GELU_TANH)])
    w_dn = [r(2, d, 4 * d, sc=0.02) for _ in range(reps)]
    gate = r(1, d)
    for i in range(reps):
        ops.gemm([ops.gemm_problem(mi, w_dn[i][0], b_up[0, :d], xi, ops.EPI_GATED_RESIDUAL, xi, gate),
                  ops.gemm_problem(mt, w_dn[i][1], b_up[1, :d], xt, ops.EPI_GATED_RESIDUAL, xt, gate)])
if which in ("all", "gemm", "qkv"):
    from vllm_omni_draft_b200 import rope as _rope
    w_qkv = [r(2, 3 * d, d, sc=0.02) for _ in range(reps)]
    b_qkv = r(2, 3 * d)
    fr_ = _rope.joint_cos_sin((16, 56, 56), 10000.0, [(1, 64, 64)], s_txt).to(dev)
    nw_ = r(4, Dh) + 1
    qkv_o = torch.empty(1, S, 3 * d, device=dev, dtype=torch.bfloat16)
    kw = dict(epilogue=ops.EPI_QKV_NORM_ROPE, freqs=fr_, heads=H, head_dim=Dh)
    for i in range(reps):
        ops.gemm([ops.gemm_problem(xi, w_qkv[i][0], b_qkv[0], qkv_o[:, s_txt:], qk_norm=(nw_[0], nw_[1]), rope_pos0=s_txt, **kw),
                  ops.gemm_problem(xt, w_qkv[i][1], b_qkv[1], qkv_o[:, :s_txt], qk_norm=(nw_[2], nw_[3]), rope_pos0=0, **kw)])
if which in ("all", "attn"):
    qkv = r(1, S, 3 * d)
    q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
    o = torch.empty(1, S, H, Dh, device=dev, dtype=torch.bfloat16)
    for i in range(reps):
        ops.attention(q, k, v, o)
if which in ("all", "elem"):
    mod = r(1, 6 * d, sc=0.1)
    out = torch.empty(s_img, d, device=dev, dtype=torch.bfloat16)
    qkv = r(S, 3 * d)
    from vllm_omni_draft_b200 import rope
    fr = rope.joint_cos_sin((16, 56, 56), 10000.0, [(1, 64, 64)], s_txt).to(dev)
    nw = r(4, Dh) + 1
    lat = [r(1, s_img, 64) for _ in range(4)]
    for i in range(reps):
        ops.adaln_modulate(xi.view(s_img, d), mod[:, :d], mod[:, d:2 * d], out, s_img)
        ops.qknorm_rope(qkv, H, Dh, S, s_txt, nw[0], nw[1], nw[2], nw[3], fr)
        ops.cfg_euler_step(lat[0], lat[1], lat[2], lat[3], 4.0, dt=-0.02)
torch.cuda.synchronize()
print("ok")
