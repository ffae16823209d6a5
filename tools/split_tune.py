"""Combined launch (img + txt problem in one persistent kernel) against one launch per problem, each with its
own tuned schedule, at ONE rank's GEMM shapes under cfg2 x ulysses-P (1-GPU emulation: same shapes, local stores).
What ops.tune_gemm decides at session setup, shown with graph-timed numbers.   usage: split_tune.py P [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops  # noqa: E402
from vllm_omni_draft_b200.roofline import _time_graph  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 4
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1
d, s_img, s_txt = 3072, 4096 // P, 512 // P
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).bfloat16()
emp = lambda *s: torch.empty(*s, device=dev, dtype=torch.bfloat16)
nw, reps = 6, 24
xi, xt = r(B, s_img, d), r(B, s_txt, d)
w_up, b_up = [r(2, 4 * d, d, sc=0.02) for _ in range(nw)], r(2, 4 * d)
mi, mt = r(B, s_img, 4 * d), r(B, s_txt, 4 * d)
gate, ho, eo = torch.zeros(B, d, device=dev, dtype=torch.bfloat16), emp(B, s_img, d), emp(B, s_txt, d)
qi, qt = emp(B, s_img, 3 * d), emp(B, s_txt, 3 * d)
w_dn = [w.view(2, d, 4 * d) for w in w_up]
w_qkv = [w.view(2, 4 * d, d)[:, :3 * d].contiguous() for w in w_up[:4]]
w_o = [w.view(2, 4 * d, d)[:, :d].contiguous() for w in w_up[:4]]
cases = {
    "qkv": lambda i: [ops.gemm_problem(xi, w_qkv[i % 4][0], b_up[0, :3 * d], qi, 0),
                      ops.gemm_problem(xt, w_qkv[i % 4][1], b_up[1, :3 * d], qt, 0)],
    "out": lambda i: [ops.gemm_problem(xi, w_o[i % 4][0], b_up[0, :d], ho, 2, ho, gate),
                      ops.gemm_problem(xt, w_o[i % 4][1], b_up[1, :d], eo, 2, eo, gate)],
    "mlp_up": lambda i: [ops.gemm_problem(xi, w_up[i % nw][0], b_up[0], mi, 1),
                         ops.gemm_problem(xt, w_up[i % nw][1], b_up[1], mt, 1)],
    "mlp_down": lambda i: [ops.gemm_problem(mi, w_dn[i % nw][0], b_up[0, :d], ho, 2, ho, gate),
                           ops.gemm_problem(mt, w_dn[i % nw][1], b_up[1, :d], eo, 2, eo, gate)],
}
tot_c = tot_s = tot_pick = 0.0
for name, mk in cases.items():
    probs = mk(0)
    key = ops._gemm_key(probs)
    chosen = ops.tune_gemm(probs, allow_split=True)
    picked_split = key in ops._split_keys
    ops._split_keys.discard(key)
    tc = _time_graph(lambda i: ops.gemm(mk(i)), reps)
    ops._split_keys.add(key)
    ts = _time_graph(lambda i: ops.gemm(mk(i)), reps)
    if not picked_split:
        ops._split_keys.discard(key)
    tot_c, tot_s, tot_pick = tot_c + tc, tot_s + ts, tot_pick + (ts if picked_split else tc)
    print(f"SPLIT_TUNE P={P} B={B} {name:9s} rows {B}x({s_img}+{s_txt}): combined {tc * 1e6:7.1f} us (pair,bn)={chosen}  "
          f"split {ts * 1e6:7.1f} us  -> tuner picks {'split' if picked_split else 'combined'}")
print(f"SPLIT_TUNE P={P} B={B} sum of the four GEMMs: combined {tot_c * 1e6:.1f} us, split {tot_s * 1e6:.1f} us, "
      f"as tuned {tot_pick * 1e6:.1f} us")
