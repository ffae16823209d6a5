"""Per-kernel device time of ONE rank's share of an MMDiT block under cfg2 x ulysses-P (1-GPU emulation:
same shapes, no peer traffic): the compute floor of the multi-GPU step.  usage: rank_shapes.py P [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops, rope  # noqa: E402
from vllm_omni_draft_b200.roofline import _time_graph  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 4
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1
d, H, Dh, s_img_full, s_txt_full = 3072, 24, 128, 4096, 512
s_img, s_txt = s_img_full // P, s_txt_full // P
S_loc, S = s_img + s_txt, s_img_full + s_txt_full
hs = H // P
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).bfloat16()
emp = lambda *s: torch.empty(*s, device=dev, dtype=torch.bfloat16)
nw, reps = 8, 24
xi, xt = r(B, s_img, d), r(B, s_txt, d)
w_up, b_up = [r(2, 4 * d, d, sc=0.02) for _ in range(nw)], r(2, 4 * d)
mi, mt = emp(B, s_img, 4 * d), emp(B, s_txt, 4 * d)
gate, ho, eo = r(B, d), emp(B, s_img, d), emp(B, s_txt, d)
res = {}
t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(xi, w_up[i % nw][0], b_up[0], mi, 1),
                                    ops.gemm_problem(xt, w_up[i % nw][1], b_up[1], mt, 1)]), reps)
res["gemm_mlp_up"] = (t, 2.0 * B * S_loc * 4 * d * d)
w_dn = [w.view(2, d, 4 * d) for w in w_up]
t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(mi, w_dn[i % nw][0], b_up[0, :d], ho, 2, xi, gate),
                                    ops.gemm_problem(mt, w_dn[i % nw][1], b_up[1, :d], eo, 2, xt, gate)]), reps)
res["gemm_mlp_down"] = (t, 2.0 * B * S_loc * 4 * d * d)
w_qkv = [w.view(2, 4 * d, d)[:, :3 * d].contiguous() for w in w_up[:4]]
qkv = emp(B, S_loc, 3 * d)
t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(xi, w_qkv[i % 4][0], b_up[0, :3 * d], qkv[:, s_txt:], 0),
                                    ops.gemm_problem(xt, w_qkv[i % 4][1], b_up[1, :3 * d], qkv[:, :s_txt], 0)]), reps)
res["gemm_qkv"] = (t, 2.0 * B * S_loc * 3 * d * d)
w_o = [w.view(2, 4 * d, d)[:, :d].contiguous() for w in w_up[:4]]
back = r(P, B, S_loc, hs * Dh)
t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(back[:, :, s_txt:], w_o[i % 4][0], b_up[0, :d], ho, 2, xi, gate, a_k_chunked=True),
                                    ops.gemm_problem(back[:, :, :s_txt], w_o[i % 4][1], b_up[1, :d], eo, 2, xt, gate, a_k_chunked=True)]), reps)
res["gemm_out_proj(k-chunked A)"] = (t, 2.0 * B * S_loc * d * d)
full = r(B, S, 3, hs, Dh)
o = emp(B, S, hs, Dh)
t = _time_graph(lambda i: ops.attention(full[:, :, 0], full[:, :, 1], full[:, :, 2], o), reps)
res["attention"] = (t, 4.0 * B * S * S * hs * Dh)
mod = r(1, 6 * d, sc=0.1)
xo_i, xo_t = emp(B * s_img, d), emp(B * s_txt, d)
sh, sc = mod[:, :d].expand(B, d), mod[:, d:2 * d].expand(B, d)
t = _time_graph(lambda i: ops.adaln_modulate_pair([(xi.view(-1, d), sh, sc, xo_i, s_img), (xt.view(-1, d), sh, sc, xo_t, s_txt)]), reps)
res["adaln_pair (x2 per block)"] = (t, 0)
fr = rope.joint_cos_sin((16, 56, 56), 10000.0, [(1, 64, 64)], s_txt_full)[:S_loc].contiguous().to(dev)
nwt = r(4, Dh) + 1
t = _time_graph(lambda i: ops.qknorm_rope(qkv.view(B * S_loc, 3 * d), H, Dh, S_loc, s_txt, nwt[0], nwt[1], nwt[2], nwt[3], fr), reps)
res["qknorm_rope (in place, no scatter)"] = (t, 0)
tot = 0.0
for k, (t, fl) in res.items():
    mult = 2 if k.startswith("adaln") else 1
    tot += t * mult
    print(f"{k:40s} {t * 1e6:8.1f} us" + (f"  {fl / t / 1e12:7.1f} TFLOP/s" if fl else ""))
print(f"block sum {tot * 1e6:.1f} us -> x60 = {tot * 60 * 1e3:.2f} ms per forward (P={P}, B={B})")
