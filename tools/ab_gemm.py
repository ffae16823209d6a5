#!/usr/bin/env python
"""Interleaved A/B of GEMM knobs on ONE box in ONE process (boxes and even successive runs on a box differ by several
percent under the power cap, so only interleaved medians mean anything).  For each of the block's four GEMM launches
(cfg-2 shapes, batch 2, fp32 residual streams) and each candidate setting, the 24-launch graph is timed `rounds`
times, alternating settings inside every round.   usage: ab_gemm.py group_m 0 6 9 12 | ab_gemm.py pdl 0 1"""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from vllm_omni_draft_b200 import _lib, ops, rope  # noqa: E402


def main():
    knob = sys.argv[1] if len(sys.argv) > 1 else "group_m"
    vals = [int(v) for v in sys.argv[2:]] or [0, 9]
    rounds, reps, nw = 7, 24, 6
    lib = _lib.lib()
    lib.qi_gemm_set_debug.argtypes, lib.qi_gemm_set_debug.restype = [_lib.C.c_int], _lib.C.c_int
    setter = {"group_m": lib.qi_gemm_set_group_m, "pdl": lib.qi_set_pdl, "dbg": lib.qi_gemm_set_debug}[knob]
    d, H, Dh, s_img, s_txt, B = 3072, 24, 128, 4096, 512, 2
    S = s_img + s_txt
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).bfloat16()
    r32 = lambda *s, sc=1.0: torch.randn(*s, device=dev, generator=g) * sc
    emp = lambda *s: torch.empty(*s, device=dev, dtype=torch.bfloat16)
    xi, xt = r(B, s_img, d), r(B, s_txt, d)
    w = [r(2, 4 * d, d, sc=0.02) for _ in range(nw)]
    bias = r(2, 4 * d)
    mi, mt = emp(B, s_img, 4 * d), emp(B, s_txt, 4 * d)
    gate, h32, e32 = r32(1, d, sc=0.1).expand(B, d), r32(B, s_img, d), r32(B, s_txt, d)
    att, qkv = r(B, S, d), emp(B, S, 3 * d)
    fr = rope.joint_cos_sin((16, 56, 56), 10000.0, [(1, 64, 64)], s_txt).to(dev)
    nwt = r(4, Dh) + 1
    qk = dict(epilogue=ops.EPI_QKV_NORM_ROPE, freqs=fr, heads=H, head_dim=Dh)
    wd = [x.view(2, d, 4 * d) for x in w]
    wq = [x[:, :3 * d].contiguous() for x in w[:4]]
    wo = [x[:, :d].contiguous() for x in w[:4]]
    launches = {
        "mlp_up": (lambda i: ops.gemm([ops.gemm_problem(xi, w[i % nw][0], bias[0], mi, 1), ops.gemm_problem(xt, w[i % nw][1], bias[1], mt, 1)]), 2.0 * B * S * 4 * d * d),
        "mlp_down": (lambda i: ops.gemm([ops.gemm_problem(mi, wd[i % nw][0], bias[0, :d], h32, 2, h32, gate), ops.gemm_problem(mt, wd[i % nw][1], bias[1, :d], e32, 2, e32, gate)]), 2.0 * B * S * 4 * d * d),
        "qkv": (lambda i: ops.gemm([ops.gemm_problem(xi, wq[i % 4][0], bias[0, :3 * d], qkv[:, s_txt:], qk_norm=(nwt[0], nwt[1]), rope_pos0=s_txt, **qk),
                                    ops.gemm_problem(xt, wq[i % 4][1], bias[1, :3 * d], qkv[:, :s_txt], qk_norm=(nwt[2], nwt[3]), rope_pos0=0, **qk)]), 2.0 * B * S * 3 * d * d),
        "out_proj": (lambda i: ops.gemm([ops.gemm_problem(att[:, s_txt:], wo[i % 4][0], bias[0, :d], h32, 2, h32, gate), ops.gemm_problem(att[:, :s_txt], wo[i % 4][1], bias[1, :d], e32, 2, e32, gate)]), 2.0 * B * S * d * d),
    }
    graphs = {}
    for name, (fn, _) in launches.items():
        for v in vals:
            setter(v)
            for i in range(2):
                fn(i)
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                for i in range(reps):
                    fn(i)
            graphs[(name, v)] = gr
    setter({"group_m": -1, "pdl": 1, "dbg": 0}[knob])
    times = {k: [] for k in graphs}
    for _ in range(rounds):
        for k, gr in graphs.items():
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            gr.replay()
            e1.record()
            torch.cuda.synchronize()
            times[k].append(e0.elapsed_time(e1) / reps * 1e3)
    out = {}
    for (name, v), ts in times.items():
        us = statistics.median(ts)
        out.setdefault(name, {})[f"{knob}={v}"] = {"us": round(us, 1), "tflops": round(launches[name][1] / us / 1e6, 0)}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
