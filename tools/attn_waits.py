"""Where do the attention kernel's MMA-issuer and softmax warps wait?  Cycle counters (qi_attention_set_debug, a
tuning aid outside the ABI) summed over the CTAs of one launch at the cfg-2 shape, printed per CTA.

Needs the experiments library for any variant other than 0, 3, 8, 11 (and for the cycle counters / DIAG modes):
    make -C tools/experiments && QI_LIB_PATH=tools/experiments/libqi_b200_exp.so python tools/attn_waits.py ...
"""
import ctypes as C
import os
import sys

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
buf = torch.zeros(16, device="cuda", dtype=torch.int64)
names = {0: "mma total", 1: "mma wait v_full", 2: "mma wait p_half", 3: "mma wait p_full", 4: "mma wait k_full",
         5: "mma wait q+k first", 8: "softmax total", 9: "softmax wait s_full", 10: "softmax wait o_done",
         11: "softmax S load", 12: "softmax row max", 13: "softmax O rescale", 14: "softmax P first half",
         15: "softmax P second half"}
for var in [int(x) for x in (sys.argv[1:] or ["11", "19"])]:
    n_cta = ((S + 127) // 128 if var >= 40 else (S + 255) // 256) * H * B   # variants 40+: one Q tile per CTA
    ops.attention_set_variant(var)
    for _ in range(3):
        ops.attention(q, k, v, o)
    torch.cuda.synchronize()
    buf.zero_()
    lib.qi_attention_set_debug(buf.data_ptr())
    ops.attention(q, k, v, o)
    torch.cuda.synchronize()
    lib.qi_attention_set_debug(None)
    c = buf.tolist()
    print(f"variant {var}: per CTA (of {n_cta}), {(S + 127) // 128} K/V tiles")
    for i, n in names.items():
        base = c[0] if i < 8 else c[8]
        print(f"  {n:22s} {c[i] / n_cta:10.0f} clk  {100.0 * c[i] / max(base, 1):5.1f} %")
ops.attention_set_variant(-1)
