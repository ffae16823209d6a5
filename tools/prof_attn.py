"""Launches the attention kernel alone at the cfg-2 shape for ncu (variant from QI_ATTN_VARIANT)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops  # noqa: E402
d, H, Dh, S = 3072, 24, 128, 4608
g = torch.Generator(device="cuda").manual_seed(0)
qkv = torch.randn(1, S, 3 * d, device="cuda", generator=g).bfloat16()
q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
o = torch.empty(1, S, H, Dh, device="cuda", dtype=torch.bfloat16)
for i in range(3):
    ops.attention(q, k, v, o)
torch.cuda.synchronize()
print("ok")
