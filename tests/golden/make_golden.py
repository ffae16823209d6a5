"""Generates tests/golden/*.pt from the oracle (the reference ships no golden vectors).
Run from the repo root:  python tests/golden/make_golden.py"""
import hashlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import qwen_image_oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def sha(t: torch.Tensor) -> str:
    return hashlib.sha256(t.contiguous().numpy().tobytes()).hexdigest()


def main():
    torch.manual_seed(0)
    cfg = O.tiny_config()
    sd = O.init_state_dict(cfg, seed=0, bias_std=0.02, norm_weight_std=0.1)
    lat, cond, uncond, mask, shapes = O.synthetic_inputs(cfg, 256, 256, 32)
    fwd = O.transformer_forward(sd, cfg, lat, cond, mask, torch.tensor([0.7446]).bfloat16().float(), shapes, [32])
    final, trace = O.denoise_loop(sd, cfg, lat, cond, mask, uncond, mask, 4, 4.0, img_shapes=shapes, return_all=True)
    ts, sig = O.flowmatch_sigmas(4, 256)
    ts50, sig50 = O.flowmatch_sigmas(50, 4096)
    torch.save({"forward_sample": fwd, "step_latents": torch.stack(trace), "final": final,
                "timesteps4": ts, "sigmas4": sig, "timesteps50_4096": ts50, "sigmas50_4096": sig50},
               os.path.join(HERE, "tiny_cfg1.pt"))
    # integer tables
    tables = {}
    for hl in (32, 128, 166):
        src = torch.arange(16 * hl * hl, dtype=torch.int64).view(1, 16, hl, hl)
        idx = O.pack_latents(src).reshape(-1)
        tables[f"pack_sha_{hl}"] = sha(idx)
        if hl == 32:
            tables["pack_idx_32"] = idx.to(torch.int32)
    for (f, h, w), txt in (((1, 16, 16), 32), ((1, 64, 64), 512), ((1, 83, 83), 512), ((1, 128, 128), 512)):
        ip, tp = O.rope_position_ids([(f, h, w)], txt)
        tables[f"rope_sha_{h}x{w}_{txt}"] = sha(torch.cat([tp, ip]).to(torch.int32))
        if h == 16:
            tables["rope_pos_16x16_32"] = torch.cat([tp, ip]).to(torch.int32)
    ip, tp = O.rope_position_ids([(1, 8, 8), (1, 6, 10)], 7)       # edit-style: two images
    tables["rope_pos_two_images"] = torch.cat([tp, ip]).to(torch.int32)
    torch.save(tables, os.path.join(HERE, "int_tables.pt"))
    print("wrote", os.listdir(HERE))


if __name__ == "__main__":
    main()
