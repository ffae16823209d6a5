"""Parity of the CUDA path at the BENCHMARKED depth (helper for tests/test_depth_gpu.py; also a script).

bench.py times 60 MMDiT blocks and a 50-step schedule (BASELINE.json configs[1]); this module compares exactly
that model with the oracle executed in **fp32 on the GPU** on the same weights:

* the 20B model is random-initialised once on the device in bf16 (`init_random_`, with non-zero biases and
  non-trivial norm weights); the oracle reads the SAME bf16-rounded values, upcast to fp32 tensor by tensor
  through `LazyStateDict` (a 41 GB model, never a second 82 GB copy);
* `forward_depth` logs the relative-L2 error of both residual streams after every block and of `.sample`;
* `loop_depth` runs N true-CFG FlowMatch steps through `B200QwenImagePipeline.__call__` and through the oracle's
  operators, compares the per-step latents, and decodes both final latents with the product's VAE decoder at the
  real AutoencoderKLQwenImage dimensions for the PSNR gate.

The loop replaced is /root/reference/src/qwen_image/runner_adapter.py:127-141 and
src/performance_enhancement/test_performance.py:277 at full depth.
"""
from __future__ import annotations

import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import qwen_image_oracle as O  # noqa: E402


LazyStateDict = O.ParamView      # diffusers-named view of the CUDA model's fused parameters, upcast on access


def build_model(num_layers: int = 60, seed: int = 5, device: str = "cuda:0"):
    from vllm_omni_draft_b200.config import MMDiTConfig
    from vllm_omni_draft_b200.transformer import B200QwenImageTransformer2DModel
    tr = B200QwenImageTransformer2DModel(MMDiTConfig(num_layers=num_layers), device=device)
    tr.init_random_(seed, bias_std=0.02, norm_weight_std=0.1)
    return tr, O.OracleConfig(num_layers=num_layers)


def _inputs(B: int, hh: int, ww: int, s_txt: int, device: str, seed: int = 11):
    g = torch.Generator().manual_seed(seed)
    lat = torch.randn(1, hh * ww, 64, generator=g).expand(B, -1, -1).contiguous().to(device, torch.bfloat16)
    enc = torch.randn(B, s_txt, 3584, generator=g).to(device, torch.bfloat16)
    mask = torch.ones(B, s_txt, dtype=torch.bool, device=device)
    return lat, enc, mask


@torch.no_grad()
def forward_depth(tr, ocfg, B: int = 2, hh: int = 64, ww: int = 64, s_txt: int = 512, t_value: float = 0.625,
                  device: str = "cuda:0") -> dict:
    """One forward at full depth: CUDA path vs fp32 oracle, error after every block."""
    lat, enc, mask = _inputs(B, hh, ww, s_txt, device)
    t = torch.full((B,), t_value, device=device)            # bf16-representable: both sides see the same t
    shapes = [[(1, hh, ww)]] * B
    ours_h, ours_e = [], []

    def hook(l, h, e):
        ours_h.append(h.clone())
        ours_e.append(e.clone())

    out = tr(hidden_states=lat, encoder_hidden_states=enc, encoder_hidden_states_mask=mask, timestep=t,
             img_shapes=shapes, txt_seq_lens=[s_txt] * B, return_dict=False, _block_hook=hook)[0]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ref, inter, _ = O.transformer_forward(LazyStateDict(tr), ocfg, lat.float(), enc.float(), mask, t, shapes,
                                          [s_txt] * B, return_intermediates=True)
    torch.cuda.synchronize()
    oracle_s = time.perf_counter() - t0
    per_block = []
    for l, (rh, re) in enumerate(inter):
        per_block.append({"block": l, "h": O.rel_l2(ours_h[l].float(), rh), "e": O.rel_l2(ours_e[l].float(), re)})
    err = O.rel_l2(out.float(), ref)
    return {"what": f"{ocfg.num_layers}-block forward, B={B}, S_img={hh * ww}, S_txt={s_txt}, CUDA path vs fp32 oracle on GPU",
            "rel_l2_sample": err, "per_block": per_block, "oracle_fp32_seconds": oracle_s,
            "h_rms": float(inter[-1][0].pow(2).mean().sqrt()), "sample_rms": float(ref.pow(2).mean().sqrt())}


@torch.no_grad()
def loop_depth(tr, ocfg, steps: int = 8, height: int = 1024, width: int = 1024,
               s_txt: int = 512, device: str = "cuda:0", use_cuda_graph: bool = True, vae=None) -> dict:
    """A complete `steps`-step true-CFG FlowMatch denoise at full depth: per-step latents of the
    pipeline call vs the oracle's operators in fp32 (same bf16-rounded timesteps upstream's bf16 loop feeds), and
    the PSNR of the two decoded images (same decoder on both sides, real VAE dimensions)."""
    from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
    Hl, Wl = O.latent_hw(height, width)
    hh, ww = Hl // 2, Wl // 2
    g = torch.Generator().manual_seed(42)
    lat = O.pack_latents(torch.randn(1, 1, 16, Hl, Wl, generator=g)).to(device, torch.bfloat16)
    cond = torch.randn(1, s_txt, 3584, generator=torch.Generator().manual_seed(1)).to(device, torch.bfloat16)
    uncond = torch.randn(1, s_txt, 3584, generator=torch.Generator().manual_seed(2)).to(device, torch.bfloat16)
    mask = torch.ones(1, s_txt, dtype=torch.bool, device=device)
    shapes = [[(1, hh, ww)]]
    pipe = B200QwenImagePipeline(tr, vae=vae)
    pipe.use_cuda_graph = use_cuda_graph
    final = pipe(prompt_embeds=cond, prompt_embeds_mask=mask, negative_prompt_embeds=uncond,
                 negative_prompt_embeds_mask=mask, true_cfg_scale=4.0, num_inference_steps=steps, height=height, width=width,
                 latents=lat, output_type="latent", record_steps=True).images
    ours = [x.clone() for x in pipe.last_step_latents]
    torch.cuda.synchronize()
    # oracle loop, fp32 arithmetic on the GPU (a complete `steps`-step schedule: sigma 1 -> 0.02 -> 0)
    timesteps, sigmas = O.flowmatch_sigmas(steps, hh * ww)
    sd = LazyStateDict(tr)
    x = lat.float()
    per_step = []
    t0 = time.perf_counter()
    for i, t in enumerate(timesteps):
        t_in = (t.to(torch.bfloat16) / 1000).float().expand(1).to(device)     # upstream: t.to(bf16) / 1000
        c = O.transformer_forward(sd, ocfg, x, cond.float(), mask, t_in, shapes, [s_txt])
        u = O.transformer_forward(sd, ocfg, x, uncond.float(), mask, t_in, shapes, [s_txt])
        v = O.true_cfg_combine(c, u, 4.0)
        x = O.scheduler_step(v, x, sigmas[i].to(device), sigmas[i + 1].to(device))
        per_step.append({"step": i, "rel_l2": O.rel_l2(ours[i].float(), x)})
    torch.cuda.synchronize()
    res = {"what": f"{steps}-step true-CFG denoise, {ocfg.num_layers} blocks, "
                   f"{height}x{width}, S_txt={s_txt}, pipeline call (graph={use_cuda_graph}) vs fp32 oracle on GPU",
           "per_step": per_step, "oracle_fp32_seconds": time.perf_counter() - t0,
           "rel_l2_final": O.rel_l2(final.float(), x)}
    if vae is not None:
        from oracle import vae_oracle as V
        img_ours = pipe.decode_latents(final, Hl, Wl).float()
        img_ref = pipe.decode_latents(x.to(torch.bfloat16), Hl, Wl).float()
        res["psnr_db"] = V.psnr(img_ours.cpu(), img_ref.cpu())
        res["image_std"] = float(img_ref.std())
        res["image_shape"] = list(img_ref.shape)
    return res


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=60)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "depth_parity.json"))
    ap.add_argument("--no-loop", action="store_true")
    args = ap.parse_args()
    from vllm_omni_draft_b200.vae import B200QwenImageVAEDecoder
    tr, ocfg = build_model(args.layers)
    res = {"forward": forward_depth(tr, ocfg)}
    print("DEPTH_FORWARD", json.dumps({k: v for k, v in res["forward"].items() if k != "per_block"}))
    for r in res["forward"]["per_block"][::6] + res["forward"]["per_block"][-1:]:
        print("  block %2d  h %.3e  e %.3e" % (r["block"], r["h"], r["e"]))
    if not args.no_loop:
        vae = B200QwenImageVAEDecoder(device="cuda:0")
        _init_vae_(vae)
        res["loop"] = loop_depth(tr, ocfg, steps=args.steps, vae=vae)
        print("DEPTH_LOOP", json.dumps(res["loop"]))
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(res, f, indent=1)


def _init_vae_(vae, seed: int = 0):
    """Random-init of the real-size decoder on the device (same N(0, .05) / gamma 1 +- .1 the VAE oracle uses)."""
    g = torch.Generator(device=vae.device).manual_seed(seed)
    for n, p in vae.named_parameters():
        r = torch.randn(p.shape, device=vae.device, generator=g)
        if n.endswith("gamma"):
            p.copy_((1 + 0.1 * r).to(p.dtype))
        elif n.endswith("bias"):
            p.copy_((0.02 * r).to(p.dtype))
        else:
            p.copy_((0.05 * r).to(p.dtype))


if __name__ == "__main__":
    main()
