"""Generates tests/golden/*.pt from the oracle (the reference ships no golden vectors).
Run from the repo root:  python tests/golden/make_golden.py"""
import hashlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import qwen_image_oracle as O  # noqa: E402
from oracle import vae_oracle as V  # noqa: E402

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
    widen()


def edit_inputs(cfg):
    """tiny edit-mode case: 256x256 target (16x16 noise tokens) + a 12x20-token reference image."""
    lat, cond, uncond, mask, _ = O.synthetic_inputs(cfg, 256, 256, 32)
    g = torch.Generator().manual_seed(43)
    ref_lat = torch.randn(1, 16, 1, 24, 40, generator=g)
    shapes = [[(1, 16, 16), (1, 12, 20)]]
    return lat, ref_lat, cond, uncond, mask, shapes


def widen():
    """Fixtures for the rows built after a-e: edit loop (f2), the reference's own placeholder loops
    (a1, a2), one-frame VAE decode (f1)."""
    cfg = O.tiny_config()
    sd = O.init_state_dict(cfg, seed=0, bias_std=0.02, norm_weight_std=0.1)
    lat, ref_lat, cond, uncond, mask, shapes = edit_inputs(cfg)
    final, trace = O.edit_denoise_loop(sd, cfg, lat, O.pack_latents(ref_lat), cond, mask, uncond, mask, 3, 4.0,
                                       img_shapes=shapes, return_all=True)
    out = {"edit_step_latents": torch.stack(trace), "edit_final": final}
    g = torch.Generator().manual_seed(5)
    lat64 = torch.randn(2, 64, 8, 12, generator=g)
    cond2 = torch.randn(2, 9, cfg.joint_attention_dim, generator=g)
    mask2 = torch.ones(2, 9, dtype=torch.bool)
    ts = (torch.linspace(1.0, 0.0, steps=3) * 1000).long()
    out["gen_loop"] = O.reference_gen_loop(sd, cfg, lat64, cond2, mask2, ts, step_size=0.1,
                                           img_shapes=[(1, 8, 12)])
    out["adapter_loop"] = O.reference_adapter_loop(sd, cfg, lat64, cond2, mask2, 3, img_shapes=[(1, 8, 12)],
                                                   txt_seq_lens=[9, 9])
    vcfg = V.tiny_vae_config()
    vsd = V.init_vae_state_dict(vcfg, seed=0)
    z = torch.randn(1, 16, 1, 6, 10, generator=torch.Generator().manual_seed(6))
    out["vae_decode"] = V.decode(vsd, vcfg, z)
    out["vae_image_from_edit_final_s4"] = V.decode_packed(vsd, vcfg, final, 32, 32)[:, :, ::4, ::4].clone()
    torch.save(out, os.path.join(HERE, "tiny_widen.pt"))
    vae_encode_golden()
    print("wrote", os.listdir(HERE))


def vae_encode_golden():
    """The encoder half (row b2: pixel images in), its own small file so the older fixtures stay byte-identical."""
    vcfg = V.tiny_vae_config()
    vsd = V.init_vae_state_dict(vcfg, seed=0)
    x = torch.rand(2, 3, 1, 48, 80, generator=torch.Generator().manual_seed(8)) * 2 - 1
    out = {"vae_encode_moments": V.encode(vsd, vcfg, x),
           "vae_encode_normalized_mode": V.normalize_latents(V.encode_mode(vsd, vcfg, x), vcfg)}
    torch.save(out, os.path.join(HERE, "tiny_vae_encode.pt"))


if __name__ == "__main__":
    if "--vae-encode-only" in sys.argv:
        vae_encode_golden()
    else:
        main()
