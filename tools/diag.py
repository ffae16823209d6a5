"""GPU diagnostic: run small GEMM / attention cases, print error structure, and save the tensors
under gpurun_out/ so that a wrong descriptor/layout can be analysed offline."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops  # noqa: E402

os.makedirs("gpurun_out", exist_ok=True)
which = sys.argv[1] if len(sys.argv) > 1 else "gemm"


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


if which == "gemm":
    for (M, N, K) in [(128, 128, 64), (128, 128, 128), (128, 256, 256), (256, 128, 64)]:
        g = torch.Generator().manual_seed(0)
        a = torch.randn(M, K, generator=g).bfloat16().cuda()
        w = torch.randn(N, K, generator=g).bfloat16().cuda()
        out = torch.full((M, N), 7.0, device="cuda", dtype=torch.bfloat16)
        ops.gemm([ops.gemm_problem(a, w, None, out)])
        torch.cuda.synchronize()
        ref = a.float() @ w.float().t()
        print(f"gemm M{M} N{N} K{K}: rel={rel(out.float(), ref):.4e} out[0,:4]={out[0,:4].tolist()} ref[0,:4]={ref[0,:4].tolist()}",
              flush=True)
        torch.save({"a": a.cpu(), "w": w.cpu(), "out": out.cpu()}, f"gpurun_out/diag_gemm_{M}_{N}_{K}.pt")
elif which == "attn":
    for (S, H, Dh) in [(128, 1, 128), (256, 1, 128), (384, 2, 128), (128, 1, 64), (300, 2, 64)]:
        g = torch.Generator().manual_seed(0)
        qkv = torch.randn(1, S, 3 * H * Dh, generator=g).bfloat16().cuda()
        d = H * Dh
        q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
        out = torch.full((1, S, H, Dh), 7.0, device="cuda", dtype=torch.bfloat16)
        ops.attention(q, k, v, out)
        torch.cuda.synchronize()
        ref = torch.nn.functional.scaled_dot_product_attention(
            q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2)).transpose(1, 2)
        print(f"attn S{S} H{H} Dh{Dh}: rel={rel(out.float(), ref):.4e} out={out[0,0,0,:4].tolist()} ref={ref[0,0,0,:4].tolist()}",
              flush=True)
        torch.save({"qkv": qkv.cpu(), "out": out.cpu()}, f"gpurun_out/diag_attn_{S}_{H}_{Dh}.pt")
