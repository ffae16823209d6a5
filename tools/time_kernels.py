"""CUDA-event timing of each hot kernel alone at the cfg-2 shapes (quick iteration aid)."""
import json
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
NW = 8


def timed(fn, reps=40):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


res = {}
xi, xt = r(1, s_img, d), r(1, s_txt, d)
w_up, b_up = [r(2, 4 * d, d, sc=0.02) for _ in range(NW)], r(2, 4 * d)
mi = torch.empty(1, s_img, 4 * d, device=dev, dtype=torch.bfloat16)
mt = torch.empty(1, s_txt, 4 * d, device=dev, dtype=torch.bfloat16)
t = timed(lambda i: ops.gemm([ops.gemm_problem(xi, w_up[i % NW][0], b_up[0], mi, 1),
                              ops.gemm_problem(xt, w_up[i % NW][1], b_up[1], mt, 1)]))
res["gemm_mlp_up"] = {"us": t * 1e6, "tflops": 2.0 * S * 4 * d * d / t / 1e12}
w_dn = [w.view(2, d, 4 * d) for w in w_up]
gate = r(1, d)
ho, eo = torch.empty_like(xi), torch.empty_like(xt)
t = timed(lambda i: ops.gemm([ops.gemm_problem(mi, w_dn[i % NW][0], b_up[0, :d], ho, 2, xi, gate),
                              ops.gemm_problem(mt, w_dn[i % NW][1], b_up[1, :d], eo, 2, xt, gate)]))
res["gemm_mlp_down"] = {"us": t * 1e6, "tflops": 2.0 * S * 4 * d * d / t / 1e12}
w_qkv = [w.view(2, 4 * d, d)[:, :3 * d].contiguous() for w in w_up[:4]]
qkv = torch.empty(1, S, 3 * d, device=dev, dtype=torch.bfloat16)
t = timed(lambda i: ops.gemm([ops.gemm_problem(xi, w_qkv[i % 4][0], b_up[0, :3 * d], qkv[:, s_txt:], 0),
                              ops.gemm_problem(xt, w_qkv[i % 4][1], b_up[1, :3 * d], qkv[:, :s_txt], 0)]))
res["gemm_qkv"] = {"us": t * 1e6, "tflops": 2.0 * S * 3 * d * d / t / 1e12}
w_o = [w.view(2, 4 * d, d)[:, :d].contiguous() for w in w_up[:4]]
att = r(1, S, d)
t = timed(lambda i: ops.gemm([ops.gemm_problem(att[:, s_txt:], w_o[i % 4][0], b_up[0, :d], ho, 2, xi, gate),
                              ops.gemm_problem(att[:, :s_txt], w_o[i % 4][1], b_up[1, :d], eo, 2, xt, gate)]))
res["gemm_out_proj"] = {"us": t * 1e6, "tflops": 2.0 * S * d * d / t / 1e12}
qkv.copy_(r(1, S, 3 * d))
q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
o = torch.empty(1, S, H, Dh, device=dev, dtype=torch.bfloat16)
t = timed(lambda i: ops.attention(q, k, v, o))
res["attention"] = {"us": t * 1e6, "tflops": 4.0 * S * S * d / t / 1e12}
mod = r(1, 6 * d, sc=0.1)
out = torch.empty(s_img, d, device=dev, dtype=torch.bfloat16)
t = timed(lambda i: ops.adaln_modulate(xi.view(s_img, d), mod[:, :d], mod[:, d:2 * d], out, s_img))
res["adaln_img"] = {"us": t * 1e6, "gbs": 4.0 * s_img * d / t / 1e9}
fr = rope.joint_cos_sin((16, 56, 56), 10000.0, [(1, 64, 64)], s_txt).to(dev)
nw = r(4, Dh) + 1
q2 = qkv.view(S, 3 * d)
t = timed(lambda i: ops.qknorm_rope(q2, H, Dh, S, s_txt, nw[0], nw[1], nw[2], nw[3], fr))
res["qknorm_rope"] = {"us": t * 1e6, "gbs": 8.0 * S * d / t / 1e9}
lat = [r(1, s_img, 64) for _ in range(4)]
t = timed(lambda i: ops.cfg_euler_step(lat[0], lat[1], lat[2], lat[3], 4.0, dt=-0.02))
res["cfg_euler"] = {"us": t * 1e6, "gbs": 4.0 * s_img * 64 * 2 / t / 1e9}
blk = res["gemm_mlp_up"]["us"] + res["gemm_mlp_down"]["us"] + res["gemm_qkv"]["us"] + res["gemm_out_proj"]["us"] \
    + res["attention"]["us"] + 2 * res["adaln_img"]["us"] * (S / s_img) + res["qknorm_rope"]["us"]
res["block_sum_us"] = blk
res["step_est_ms"] = blk * 120 / 1e3
print(json.dumps(res, indent=1))
