"""Full-size checks on the GPU (BASELINE.json configs[1] shapes: d=3072, 24x128 heads, D_txt=3584, 4096 image
tokens + 512 text tokens, cond+uncond stacked as batch 2; two of the 60 blocks so it runs in seconds):

* parity of the CUDA path against the oracle executed op-by-op in bf16 ON THE GPU — i.e. what eager PyTorch
  (cuBLASLt GEMMs + F.scaled_dot_product_attention), the reference's effective kernels on a B200, produces;
* the same eager run timed beside ours as the "library baseline" of BASELINE.md §5.  Reported, and asserted
  only loosely (ours must not be slower)."""
import json
import os

import pytest
import torch

from oracle import qwen_image_oracle as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _time_ms(fn, warm=2, iters=4):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


L = 2


@pytest.fixture(scope="module")
def full_width(cuda_device):
    """Two blocks at the real width on the GPU: (oracle config, bf16 state dict on the GPU, CUDA model)."""
    from vllm_omni_draft_b200.config import MMDiTConfig
    from vllm_omni_draft_b200.transformer import B200QwenImageTransformer2DModel
    ocfg = O.OracleConfig(num_layers=L)
    sd = O.init_state_dict(ocfg, seed=3, bias_std=0.02, norm_weight_std=0.1)
    tr = B200QwenImageTransformer2DModel(MMDiTConfig(num_layers=L), device="cuda:0")
    tr.load_diffusers_state_dict(sd)
    sdg = {k: v.to("cuda:0", torch.bfloat16) for k, v in sd.items()}
    del sd
    return ocfg, sdg, tr


@pytest.mark.parametrize("name,B,hh,ww,s_txt", [
    ("cfg4_2048px", 1, 128, 128, 512),       # BASELINE configs[3]: 16384 image tokens (one CFG branch)
    ("cfg5_1328px", 2, 83, 83, 512),         # configs[4]: 83x83 = 6889 tokens (odd grid, ragged tiles), batched
    ("short_prompt", 2, 64, 64, 77),         # odd text length at 1024px
])
def test_other_configs_full_width_parity(full_width, name, B, hh, ww, s_txt):
    """The other BASELINE.json configs as parity cases: real width, real sequence lengths, against the oracle
    run in eager bf16 on the GPU."""
    ocfg, sdg, tr = full_width
    g = torch.Generator().manual_seed(hh * 1000 + s_txt)
    lat = torch.randn(B, hh * ww, 64, generator=g).to("cuda:0", torch.bfloat16)
    enc = torch.randn(B, s_txt, 3584, generator=g).to("cuda:0", torch.bfloat16)
    mask = torch.ones(B, s_txt, dtype=torch.bool, device="cuda:0")
    t = torch.full((B,), 0.35, device="cuda:0")
    shapes = [[(1, hh, ww)]] * B
    with torch.no_grad():
        ref = O.transformer_forward(sdg, ocfg, lat, enc, mask, t.bfloat16(), shapes, [s_txt] * B)
        out = tr(hidden_states=lat, encoder_hidden_states=enc, encoder_hidden_states_mask=mask, timestep=t,
                 img_shapes=shapes, txt_seq_lens=[s_txt] * B, return_dict=False)[0]
    torch.cuda.synchronize()
    err = O.rel_l2(out.float().cpu(), ref.float().cpu())
    print(f"FULL_WIDTH_PARITY {name}: B={B} S_img={hh * ww} S_txt={s_txt} rel_l2_vs_eager_bf16={err:.3e}")
    assert err < 2e-2, f"{name}: rel-L2 {err}"
    tr._ws.clear()                            # release this shape's workspace before the next case


def test_full_size_blocks_parity_and_library_baseline(full_width):
    B, s_img, s_txt = 2, 4096, 512
    ocfg, sdg, tr = full_width
    g = torch.Generator().manual_seed(11)
    lat = torch.randn(1, s_img, 64, generator=g).expand(B, -1, -1).contiguous().to("cuda:0", torch.bfloat16)
    enc = torch.randn(B, s_txt, 3584, generator=g).to("cuda:0", torch.bfloat16)
    mask = torch.ones(B, s_txt, dtype=torch.bool, device="cuda:0")
    t = torch.tensor([0.62, 0.62], device="cuda:0")
    shapes = [[(1, 64, 64)]] * B

    def eager():
        return O.transformer_forward(sdg, ocfg, lat, enc, mask, t.bfloat16(), shapes, [s_txt] * B)

    def ours():
        return tr(hidden_states=lat, encoder_hidden_states=enc, encoder_hidden_states_mask=mask, timestep=t,
                  img_shapes=shapes, txt_seq_lens=[s_txt] * B, return_dict=False)[0]

    with torch.no_grad():
        ref = eager()
        out = ours()
        torch.cuda.synchronize()
        err = O.rel_l2(out.float().cpu(), ref.float().cpu())
        # both sides round to bf16 (in different places): twice the fp32-oracle gate
        assert err < 2e-2, f"rel-L2 vs eager bf16 at full cfg-2 shapes: {err}"
        tr.tune_schedules(B, s_img, s_txt)
        ms_eager = _time_ms(eager)
        ms_ours = _time_ms(ours)
    d = ocfg.inner_dim
    S = s_img + s_txt
    flop_block = 24 * d * d * S * B + 4 * S * S * d * B
    line = {"what": "2 MMDiT blocks + in/out projections, cfg-2 shapes, batch 2 (cond+uncond)",
            "rel_l2_vs_eager_bf16": err, "eager_ms": ms_eager, "ours_ms": ms_ours,
            "eager_ms_per_block": ms_eager / L, "ours_ms_per_block": ms_ours / L,
            "speedup": ms_eager / ms_ours,
            "eager_tflops": flop_block * L / ms_eager * 1e-9, "ours_tflops": flop_block * L / ms_ours * 1e-9,
            "note": "un-hoisted modulation on both sides (timestep passed per call); no CUDA graph on either"}
    print("LIBRARY_BASELINE " + json.dumps(line))
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        with open(os.path.join(out_dir, "library_baseline.json"), "w") as f:
            json.dump(line, f)
    assert ms_ours < ms_eager
