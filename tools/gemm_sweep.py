"""A/B of the GEMM schedules (single-CTA vs CTA-pair, tile width) at the block's shapes; long graphs
(>= 100 ms) so the numbers are at sustained clocks, not burst.  python tools/gemm_sweep.py [sp]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops  # noqa: E402
from vllm_omni_draft_b200.roofline import _time_graph  # noqa: E402

d = 3072
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).bfloat16()
sp = int(sys.argv[1]) if len(sys.argv) > 1 else 1
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 1
s_img, s_txt = 4096 // sp, 512 // sp
nw = 6
W = [r(2, 4 * d, d, sc=0.02) for _ in range(nw)]
bias = r(2, 4 * d)
shapes = {"qkv": (3 * d, d, 0), "out": (d, d, 2), "up": (4 * d, d, 1), "down": (d, 4 * d, 2)}
for name, (N, K, epi) in shapes.items():
    xi, xt = r(batch, s_img, K), r(batch, s_txt, K)
    oi, ot = torch.empty(batch, s_img, N, device=dev, dtype=torch.bfloat16), torch.empty(batch, s_txt, N, device=dev, dtype=torch.bfloat16)
    ri, rt = (r(batch, s_img, N), r(batch, s_txt, N)) if epi == 2 else (None, None)
    gate = r(batch, N) if epi == 2 else None
    ws = [w.view(-1)[:2 * N * K].view(2, N, K) for w in W]
    fl = 2.0 * batch * (s_img + s_txt) * N * K
    only = sys.argv[3] if len(sys.argv) > 3 else ""
    ops.gemm_set_cta_pair(-1)
    ops.gemm_set_tile_n(0)
    fn0 = lambda i: ops.gemm([ops.gemm_problem(xi, ws[i % nw][0], bias[0, :N], oi, epi, ri, gate),
                              ops.gemm_problem(xt, ws[i % nw][1], bias[1, :N], ot, epi, rt, gate)])
    t = _time_graph(fn0, max(8, int(0.05 / _time_graph(fn0, 8))))
    print(f"sp={sp} B={batch} {name:5s} AUTO           : {t * 1e6:8.1f} us  {fl / t / 1e12:7.1f} TF/s", flush=True)
    for pair in (0, 1):
        for bn in (256, 224, 192, 160, 128):
            if only == "pair256" and not (pair == 1 and bn == 256):
                continue
            ops.gemm_set_cta_pair(pair)
            ops.gemm_set_tile_n(bn)
            fn = lambda i: ops.gemm([ops.gemm_problem(xi, ws[i % nw][0], bias[0, :N], oi, epi, ri, gate),
                                     ops.gemm_problem(xt, ws[i % nw][1], bias[1, :N], ot, epi, rt, gate)])
            t_one = _time_graph(fn, 8)
            reps = max(8, int(0.05 / t_one))
            t = _time_graph(fn, reps)
            print(f"sp={sp} B={batch} {name:5s} pair={pair} bn={bn:3d}: {t * 1e6:8.1f} us  {fl / t / 1e12:7.1f} TF/s (burst {fl / t_one / 1e12:7.1f})", flush=True)
ops.gemm_set_cta_pair(-1)
ops.gemm_set_tile_n(0)
