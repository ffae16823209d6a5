"""Per-op device timeline of one eager denoising step on every rank (CUDA events around each ops.* call):
shows where a multi-GPU step spends its time, including the waits inside the peer barriers.
torchrun --nproc-per-node N tools/mgpu_trace.py"""
import collections
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import ops  # noqa: E402
from vllm_omni_draft_b200.config import MMDiTConfig  # noqa: E402
from vllm_omni_draft_b200.launcher import ParallelDenoiser, init_distributed  # noqa: E402

world, rank, local_rank = init_distributed(int(os.environ.get("WORLD_SIZE", "1")))
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
cfg = MMDiTConfig.qwen_image_20b()
layers = int(os.environ.get("QI_TRACE_LAYERS", "60"))
if layers != 60:
    cfg = MMDiTConfig(num_layers=layers)
H = W = int(os.environ.get("QI_TRACE_HW", "1024"))
s_img, s_txt = (H // 16) * (W // 16), 512
den = ParallelDenoiser(cfg, dev, world, rank, seed=0, use_cuda_graph=False)
g1, g2, g3 = (torch.Generator().manual_seed(s) for s in (1, 2, 42))
cond = torch.randn(1, s_txt, cfg.joint_attention_dim, generator=g1).bfloat16()
uncond = torch.randn(1, s_txt, cfg.joint_attention_dim, generator=g2).bfloat16()
lat = torch.randn(1, s_img, 64, generator=g3).bfloat16()
if os.environ.get("QI_TRACE_NO_CFG", "0") == "1":     # one branch only: with QI_CFG_PARALLEL=0 on P GPUs this is one
    uncond = None                                     # rank's share of the 2P-GPU cfg x ulysses-P layout
den.setup(cond, uncond, lat, H, W, 50, true_cfg_scale=4.0)
for i in range(3):
    den.step(i)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()

records = []
names = ["gemm", "attention", "attention_scatter", "adaln_modulate_pair", "adaln_modulate", "peer_barrier",
         "qknorm_rope", "qknorm_rope_scatter", "cfg_euler_step", "rmsnorm_affine"]
orig = {n: getattr(ops, n) for n in names}
counter = collections.Counter()


def wrap(name, fn):
    def inner(*a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = fn(*a, **k)
        e1.record()
        label = name
        if name == "gemm":                      # order inside a block: qkv, out, up, down
            label = "gemm_" + ["qkv", "out", "up", "down"][counter["gemm"] % 4] if counter["gemm_started"] else "gemm_in"
            if counter["gemm_started"]:
                counter["gemm"] += 1
            counter["gemm_started"] = 1
        records.append((label, e0, e1))
        return r
    return inner


for n in names:
    setattr(ops, n, wrap(n, orig[n]))
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
den.step(3)
t1.record()
torch.cuda.synchronize()
for n in names:
    setattr(ops, n, orig[n])
agg = collections.OrderedDict()
for label, e0, e1 in records:
    a = agg.setdefault(label, [0, 0.0])
    a[0] += 1
    a[1] += e0.elapsed_time(e1) * 1e3
# gaps: time between the end of one op and the start of the next
gap = sum(records[i][2].elapsed_time(records[i + 1][1]) for i in range(len(records) - 1)) * 1e3
total = t0.elapsed_time(t1) * 1e3
lines = [f"rank {rank}: gemm schedules (pair, bn): {getattr(den.sess, 'gemm_schedules', None)}",
         f"rank {rank}: traced eager step {total / 1e3:.2f} ms, {len(records)} ops, inter-op gaps {gap / 1e3:.2f} ms"]
for k, (n, us) in agg.items():
    lines.append(f"  {k:22s} n={n:4d} total {us / 1e3:7.2f} ms  avg {us / n:7.1f} us")
for r in range(world):
    if r == rank and (r == 0 or r == world - 1):
        print("\n".join(lines), flush=True)
    if world > 1:
        dist.barrier()
den.close()
if world > 1:
    dist.destroy_process_group()
