"""GPU parity of the assembled path (transformer forward, pipeline loop) against the oracle on
identical random-init weights, seeds and latents.  Gate (BASELINE.json north_star): per-step latents
within relative-L2 <= 1e-2 in bf16."""
import pytest
import torch

from oracle import qwen_image_oracle as O

pytestmark = pytest.mark.gpu

TOL = 1e-2


def _build(ocfg, sd):
    from vllm_omni_draft_b200.config import MMDiTConfig
    from vllm_omni_draft_b200.transformer import B200QwenImageTransformer2DModel
    cfg = MMDiTConfig(num_layers=ocfg.num_layers, attention_head_dim=ocfg.attention_head_dim,
                      num_attention_heads=ocfg.num_attention_heads, joint_attention_dim=ocfg.joint_attention_dim,
                      axes_dims_rope=ocfg.axes_dims_rope)
    tr = B200QwenImageTransformer2DModel(cfg, device="cuda:0")
    tr.load_diffusers_state_dict(sd)
    return tr


@pytest.fixture(scope="module")
def tiny(cuda_device):
    ocfg = O.tiny_config()
    sd = O.init_state_dict(ocfg, seed=0, bias_std=0.02, norm_weight_std=0.1)
    return ocfg, sd, _build(ocfg, sd)


def test_forward_tiny_matches_oracle(tiny):
    ocfg, sd, tr = tiny
    lat, cond, _, mask, shapes = O.synthetic_inputs(ocfg, 256, 256, 32, batch=2)
    t = torch.tensor([0.7446, 0.7446])
    ref = O.transformer_forward(sd, ocfg, lat, cond, mask, t.bfloat16().float(), shapes, [32, 32])
    out = tr(hidden_states=lat, encoder_hidden_states=cond, encoder_hidden_states_mask=mask, timestep=t,
             img_shapes=shapes, txt_seq_lens=[32, 32]).sample
    assert out.shape == ref.shape and out.dtype == torch.bfloat16
    assert O.rel_l2(out.float().cpu(), ref) < TOL


def test_forward_distinct_timesteps_and_ragged_shapes(tiny):
    """per-sample timesteps (general boundary #1) and a non-square, odd token grid + odd text length."""
    ocfg, sd, tr = tiny
    g = torch.Generator().manual_seed(7)
    B, hh, ww, s_txt = 2, 9, 13, 21
    lat = torch.randn(B, hh * ww, 64, generator=g)
    cond = torch.randn(B, s_txt, ocfg.joint_attention_dim, generator=g)
    mask = torch.ones(B, s_txt, dtype=torch.bool)
    t = torch.tensor([0.95, 0.12])
    shapes = [[(1, hh, ww)]] * B
    ref = O.transformer_forward(sd, ocfg, lat, cond, mask, t.bfloat16().float(), shapes, [s_txt] * B)
    out = tr(hidden_states=lat, encoder_hidden_states=cond, encoder_hidden_states_mask=mask, timestep=t,
             img_shapes=shapes, txt_seq_lens=[s_txt] * B, return_dict=False)[0]
    assert O.rel_l2(out.float().cpu(), ref) < TOL


def test_reference_adapter_call_signature(tiny):
    """The reference's own loop passes int64 timesteps and may omit img_shapes/txt_seq_lens
    (src/qwen_image_gen/model.py:122-129); the drop-in must accept that call."""
    ocfg, sd, tr = tiny
    lat, cond, _, mask, _ = O.synthetic_inputs(ocfg, 256, 256, 32)
    out = tr(hidden_states=lat, encoder_hidden_states=cond, encoder_hidden_states_mask=mask,
             timestep=torch.tensor([1]).long().expand(1), guidance=None, return_dict=True)
    assert out.sample.shape == (1, 256, 64) and torch.isfinite(out.sample.float()).all()
    with pytest.raises(ValueError):
        tr(hidden_states=lat[:, :250], encoder_hidden_states=cond, timestep=torch.tensor([1.0]))


@pytest.mark.parametrize("graph", [False, True])
def test_pipeline_per_step_latents(tiny, graph):
    from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
    ocfg, sd, tr = tiny
    lat, cond, uncond, mask, shapes = O.synthetic_inputs(ocfg, 256, 256, 32)
    ref, trace = O.denoise_loop(sd, ocfg, lat, cond, mask, uncond, mask, 4, 4.0, img_shapes=shapes, return_all=True)
    pipe = B200QwenImagePipeline(tr)
    pipe.use_cuda_graph = graph
    out = pipe(prompt_embeds=cond, prompt_embeds_mask=mask, negative_prompt_embeds=uncond,
               negative_prompt_embeds_mask=mask, true_cfg_scale=4.0, num_inference_steps=4, height=256, width=256,
               latents=lat, output_type="latent", record_steps=True).images
    assert len(pipe.last_step_latents) == 4
    for i, (got, want) in enumerate(zip(pipe.last_step_latents, trace)):
        err = O.rel_l2(got.float().cpu(), want)
        assert err < TOL, f"step {i}: rel-L2 {err}"
    assert O.rel_l2(out.float().cpu(), ref) < TOL
    # against the oracle run op-by-op in bf16 (what eager diffusers would produce) the gap is similar
    sdb = {k: v.bfloat16() for k, v in sd.items()}
    refb = O.denoise_loop(sdb, ocfg, lat.bfloat16(), cond.bfloat16(), mask, uncond.bfloat16(), mask, 4, 4.0,
                          img_shapes=shapes)
    assert O.rel_l2(out.float().cpu(), refb.float()) < 2 * TOL


def test_pipeline_no_cfg_and_unpacked_output(tiny):
    from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
    ocfg, sd, tr = tiny
    lat, cond, _, mask, shapes = O.synthetic_inputs(ocfg, 256, 256, 32)
    ref = O.denoise_loop(sd, ocfg, lat, cond, mask, None, None, 3, 1.0, img_shapes=shapes)
    pipe = B200QwenImagePipeline(tr)
    out = pipe(prompt_embeds=cond, prompt_embeds_mask=mask, true_cfg_scale=1.0, num_inference_steps=3,
               height=256, width=256, latents=lat, output_type="unpacked").images
    assert out.shape == (1, 16, 1, 32, 32)
    assert O.rel_l2(out.float().cpu(), O.unpack_latents(ref, 32, 32)) < TOL


def test_block_at_20b_width(cuda_device):
    """One block at the real width (d=3072, 24x128 heads, D_txt=3584) on a shortened sequence."""
    ocfg = O.OracleConfig(num_layers=1)
    sd = O.init_state_dict(ocfg, seed=3, bias_std=0.02, norm_weight_std=0.1)
    tr = _build(ocfg, sd)
    g = torch.Generator().manual_seed(11)
    hh, ww, s_txt = 20, 24, 75
    lat = torch.randn(1, hh * ww, 64, generator=g)
    cond = torch.randn(1, s_txt, 3584, generator=g)
    mask = torch.ones(1, s_txt, dtype=torch.bool)
    t = torch.tensor([0.5])
    ref = O.transformer_forward(sd, ocfg, lat, cond, mask, t, [[(1, hh, ww)]], [s_txt])
    out = tr(hidden_states=lat, encoder_hidden_states=cond, encoder_hidden_states_mask=mask, timestep=t,
             img_shapes=[[(1, hh, ww)]], txt_seq_lens=[s_txt]).sample
    assert O.rel_l2(out.float().cpu(), ref) < TOL


_MGPU_CASES = [
    # (gpus, env) — CFG-parallel at 2, CFG x Ulysses at 4/8, pure Ulysses (QI_CFG_PARALLEL=0), both exchanges,
    # ragged shards (272x240 px -> 255 image tokens, 33 prompt tokens) and batch 2
    (2, dict(QI_GRAPH="0")),
    (2, dict(QI_GRAPH="1")),
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="peer")),
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="nccl")),
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="peer", QI_PAR_H="272", QI_PAR_W="240",
             QI_PAR_TXT="33", QI_PAR_BATCH="2")),
    (4, dict(QI_GRAPH="1", QI_SP_EXCHANGE="peer")),
    (4, dict(QI_GRAPH="1", QI_SP_EXCHANGE="nccl")),
    (4, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="peer", QI_PAR_H="272", QI_PAR_W="240",
             QI_PAR_TXT="33", QI_PAR_BATCH="2")),
    (8, dict(QI_GRAPH="1", QI_SP_EXCHANGE="peer")),
]


@pytest.mark.parametrize("n,env", _MGPU_CASES)
def test_multi_gpu_parity(cuda_device, n, env):
    """CFG-parallel (n=2), CFG x Ulysses (n=4, 8) and pure-Ulysses runs must reproduce the oracle's latents."""
    import os
    import subprocess
    import sys
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    port = 29600 + n
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(root, "tools", "mgpu_parity.py")],
                       capture_output=True, text=True, timeout=600, env=dict(os.environ, **env))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MGPU_PARITY" in r.stdout
