"""torchrun entry: CFG-parallel / Ulysses denoising on N GPUs vs the oracle (tiny config).
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/mgpu_parity.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import qwen_image_oracle as O  # noqa: E402
from vllm_omni_draft_b200.config import MMDiTConfig  # noqa: E402
from vllm_omni_draft_b200.launcher import ParallelDenoiser, init_distributed  # noqa: E402


def main():
    world, rank, local_rank = init_distributed(int(os.environ.get("WORLD_SIZE", "1")))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    ocfg = O.tiny_config()
    sd = O.init_state_dict(ocfg, seed=0, bias_std=0.02, norm_weight_std=0.1)
    # shape knobs: QI_PAR_H/W (pixels), QI_PAR_TXT (prompt tokens), QI_PAR_BATCH; 272x240 with 33 prompt
    # tokens gives ragged shards (255 image tokens)
    env = lambda k, v: int(os.environ.get(k, v))
    Hpx, Wpx, s_txt, batch = env("QI_PAR_H", 256), env("QI_PAR_W", 256), env("QI_PAR_TXT", 32), env("QI_PAR_BATCH", 1)
    lat, cond, uncond, mask, shapes = O.synthetic_inputs(ocfg, Hpx, Wpx, s_txt, batch=batch)
    steps = 4
    edit = os.environ.get("QI_PAR_EDIT", "0") == "1"     # edit mode: a 12x20-token reference image behind the noise tokens
    ref_tokens = ref_hw = None
    if edit:
        gref = torch.Generator().manual_seed(43)
        ref_tokens = O.pack_latents(torch.randn(batch, 16, 1, 24, 40, generator=gref))
        ref_hw = (12, 20)
        shapes = [[shapes[0][0], (1, 12, 20)]] * batch
    den = ParallelDenoiser(MMDiTConfig.tiny(), dev, world, rank, use_cuda_graph=os.environ.get("QI_GRAPH", "0") == "1", state_dict=sd)
    # QI_PAR_SIGMAS=1: the same schedule passed explicitly through `sigmas=` (upstream's default list) — the
    # custom-schedule plumbing of the launcher must reproduce the default run
    import numpy as np
    sig = np.linspace(1.0, 1.0 / steps, steps).tolist() if os.environ.get("QI_PAR_SIGMAS", "0") == "1" else None
    den.setup(cond.bfloat16(), uncond.bfloat16(), lat.bfloat16(), Hpx, Wpx, steps, true_cfg_scale=4.0,
              ref_h=ref_tokens.bfloat16() if edit else None, ref_hw=ref_hw, sigmas=sig)
    for i in range(steps):
        den.step(i)
    out = torch.zeros(batch, lat.shape[1], 64, dtype=torch.bfloat16).pin_memory()
    den.read_latents(out)
    torch.cuda.synchronize()
    full = out.to(dev)
    if world > 1:
        # every rank filled only its own token shard; sum over the SP ranks of the cond branch
        if den.layout.branch != 0:
            full.zero_()
        dist.all_reduce(full)
    graph_on = den.graph is not None
    if den.exchange_timed_out():
        print(f"rank {rank}: peer barrier timed out", flush=True)
        sys.exit(3)
    exch = den.exchange if den.layout.sp_degree > 1 else "-"
    den.close()
    if rank == 0:
        if edit:
            ref = O.edit_denoise_loop(sd, ocfg, lat, ref_tokens, cond, mask, uncond, mask, steps, 4.0, img_shapes=shapes)
        else:
            ref = O.denoise_loop(sd, ocfg, lat, cond, mask, uncond, mask, steps, 4.0, img_shapes=shapes)
        err = O.rel_l2(full.float().cpu(), ref)
        lay = den.layout
        print(f"MGPU_PARITY world={world} cfg={lay.cfg_degree} sp={lay.sp_degree} exchange={exch} graph={graph_on} "
              f"px={Hpx}x{Wpx} txt={s_txt} B={batch} edit={int(edit)} rel_l2={err:.4e}", flush=True)
        if not err < 1e-2:
            sys.exit(1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
