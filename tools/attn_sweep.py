"""Times every softmax variant of the attention kernel (EMU + 8*SPLIT) at the cfg-2 step shape (batch 2 =
stacked cond/uncond, 4608 tokens, 24 heads) and at the per-rank shape of cfg2 x ulysses4 (6 heads).
Variants are interleaved over several rounds (boost clocks decay over the first seconds of load, so a
sequential sweep favours whatever runs first); SM clock and board power are sampled through NVML.

Needs the experiments library for any variant other than 0, 3, 8, 11 (and for the cycle counters / DIAG modes):
    make -C tools/experiments && QI_LIB_PATH=tools/experiments/libqi_b200_exp.so python tools/attn_sweep.py ...
"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops  # noqa: E402

try:
    import pynvml
    pynvml.nvmlInit()
    _h = pynvml.nvmlDeviceGetHandleByIndex(0)

    def clk_pwr():
        return (pynvml.nvmlDeviceGetClockInfo(_h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(_h) / 1e3)
except Exception:  # pragma: no cover
    def clk_pwr():
        return (0, 0.0)

dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
variants = [int(v) for v in os.environ.get("QI_SWEEP_VARIANTS", "0,2,3,4,8,10,11,12").split(",")]
rounds = int(os.environ.get("QI_SWEEP_ROUNDS", "3"))
for (B, S, H) in [(2, 4608, 24), (1, 4608, 6)]:
    d = H * 128
    qkv = (torch.randn(B, S, 3 * d, device=dev, generator=g)).bfloat16()
    q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, 128)) for i in range(3)]
    o = torch.empty(B, S, H, 128, device=dev, dtype=torch.bfloat16)
    fl = 4.0 * B * S * S * d
    graphs = {}
    reps = 24
    for var in variants:
        ops.attention_set_variant(var)
        for _ in range(2):
            ops.attention(q, k, v, o)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for _ in range(reps):
                ops.attention(q, k, v, o)
        graphs[var] = gr
    ops.attention_set_variant(-1)
    res = {var: [] for var in variants}
    for r in range(rounds):
        for var in variants:
            gr = graphs[var]
            n_rep = int(os.environ.get("QI_SWEEP_REPLAYS", "12"))
            clks, pws = [], []
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n_rep):
                gr.replay()
            e1.record()
            while not e1.query():          # sample while the GPU works through the queued replays
                c, p = clk_pwr()
                clks.append(c)
                pws.append(p)
                time.sleep(0.005)
            torch.cuda.synchronize()
            if not clks:
                clks, pws = [0], [0.0]
            us = e0.elapsed_time(e1) * 1e3 / (n_rep * reps)
            res[var].append((us, sorted(clks)[len(clks) // 2], sorted(pws)[len(pws) // 2]))
    for var in variants:
        rr = sorted(res[var])
        us, c, p = rr[len(rr) // 2]
        print(f"B={B} S={S} H={H} variant={var:2d} (emu {var & 7}/8, split {var >> 3}): median {us:8.1f} us "
              f"{fl / us / 1e6:7.1f} TFLOP/s  sm {c} MHz {p:.0f} W   all: " + " ".join(f"{x[0]:.1f}" for x in res[var]), flush=True)
