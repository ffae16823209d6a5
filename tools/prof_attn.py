"""Launches attention alone at the cfg-2 shape (batch 2, 4608 tokens, 24 x 128 heads) for ncu: our kernel in each
variant listed in QI_ATTN_VARIANTS (default "11"), then — the library bar, VERDICT r1 item 4 — cuDNN's fused SDPA
on the same tensors, so one report holds both kernels' pipe utilisation, instruction mix and launch geometry.

Needs the experiments library for any variant other than 0, 3, 8, 11 (and for the cycle counters / DIAG modes):
    make -C tools/experiments && QI_LIB_PATH=tools/experiments/libqi_b200_exp.so python tools/prof_attn.py ...
"""
import os
import sys

import torch
import torch.nn.functional as F
from torch.nn.attention import SDPBackend, sdpa_kernel

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops  # noqa: E402

d, H, Dh, S, B = 3072, 24, 128, 4608, 2
g = torch.Generator(device="cuda").manual_seed(0)
qkv = torch.randn(B, S, 3 * d, device="cuda", generator=g).bfloat16()
q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
o = torch.empty(B, S, H, Dh, device="cuda", dtype=torch.bfloat16)
for var in [int(x) for x in os.environ.get("QI_ATTN_VARIANTS", "11").split(",")]:
    ops.attention_set_variant(var)
    for i in range(2):
        ops.attention(q, k, v, o)
ops.attention_set_variant(-1)
torch.cuda.synchronize()
if os.environ.get("QI_PROF_SDPA", "1") == "1":
    qt, kt, vt = (x.transpose(1, 2) for x in (q, k, v))
    with sdpa_kernel([SDPBackend.CUDNN_ATTENTION]):
        for i in range(2):
            F.scaled_dot_product_attention(qt, kt, vt)
    torch.cuda.synchronize()
print("ok")
