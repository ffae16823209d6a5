"""Timing experiments on the Q-in-TMEM attention kernel (variants 40+): QI_ATTN_TS_DIAG = 0 (real), 1 (softmax only
hands the barriers on: the MMA/TMA pipeline alone), 2 (S load + P store without the math), 3 (as 1, and the P*V issuer
does not wait for P) — read per process, so this script re-executes itself per mode.  Prints us per launch at the
cfg-2 step shape, the softmax phase counters of one thread and (QI_DIAG_TRACE=1) the timeline of CTA (0,0,0):
K/V TMA issue, K ready (Q*K^T issuer), P ready / step issued (P*V issuer), S ready / P published (softmax).

Needs the experiments library for any variant other than 0, 3, 8, 11 (and for the cycle counters / DIAG modes):
    make -C tools/experiments && QI_LIB_PATH=tools/experiments/libqi_b200_exp.so python tools/attn_ts_diag.py ...
"""
import ctypes as C
import os
import subprocess
import sys

if len(sys.argv) < 2:
    for mode in os.environ.get("QI_DIAG_MODES", "0,1,2,3").split(","):
        env = dict(os.environ, QI_ATTN_TS_DIAG=mode)
        subprocess.run([sys.executable, __file__, mode], env=env, check=False)
    sys.exit(0)

import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import _lib, ops  # noqa: E402

lib = _lib.lib()
lib.qi_attention_set_debug.argtypes, lib.qi_attention_set_debug.restype = [C.c_void_p], C.c_int
d, H, Dh, S, B = 3072, 24, 128, 4608, 2
g = torch.Generator(device="cuda").manual_seed(0)
qkv = torch.randn(B, S, 3 * d, device="cuda", generator=g).bfloat16()
q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
o = torch.empty(B, S, H, Dh, device="cuda", dtype=torch.bfloat16)
buf = torch.zeros(16 + 8 * 96, device="cuda", dtype=torch.int64)
for var in [int(x) for x in os.environ.get("QI_DIAG_VARIANTS", "43,59").split(",")]:
    ops.attention_set_variant(var)
    for _ in range(3):
        ops.attention(q, k, v, o)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        ops.attention(q, k, v, o)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    buf.zero_()
    lib.qi_attention_set_debug(buf.data_ptr())
    ops.attention(q, k, v, o)
    torch.cuda.synchronize()
    lib.qi_attention_set_debug(None)
    c = buf.tolist()
    n_cta = ((S + 127) // 128) * H * B
    n_half = (S + 63) // 64
    print(f"diag {sys.argv[1]} variant {var}: {us:7.1f} us  | P*V issuer: {c[0] / n_cta / n_half:6.0f} clk per half-step | "
          f"softmax wg0 total {c[8] / n_cta:7.0f}, wait s {c[9] / n_cta:7.0f} per CTA; per own half: "
          f"S load {c[11] / n_cta / (n_half / 2):5.0f}, max {c[12] / n_cta / (n_half / 2):5.0f}, rescale {c[13] / n_cta / (n_half / 2):5.0f}, "
          f"P {c[14] / n_cta / (n_half / 2):5.0f}, publish {c[15] / n_cta / (n_half / 2):5.0f}", flush=True)
    if os.environ.get("QI_DIAG_TRACE", "1") == "1":
        t = [c[16 + 8 * h:16 + 8 * h + 8] for h in range(min(n_half, 96))]
        t0 = min(x for row in t for x in row if x > 0)
        print("    timeline of CTA (0,0,0), clk since the first event (-1 = not recorded): half | K tma issued, V tma issued | "
              "K ready, -, P ready, P*V issued | softmax: S ready, P published")
        for h in list(range(0, 20)) + list(range(40, 52)):
            print(f"    {h:3d} | " + " ".join(f"{(x - t0) if x > 0 else -1:7d}" for x in t[h]))
ops.attention_set_variant(-1)
