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


# ------------------------------------------------------------------ rows built after a-e
def test_forward_two_images_and_out_rows(tiny):
    """Edit-style forward: two entries in img_shapes (RoPE frame index = image index); every
    row of the prediction is compared, so the reference-image rows' own RoPE / projections are checked;
    `_out_rows` must return exactly the leading rows."""
    ocfg, sd, tr = tiny
    g = torch.Generator().manual_seed(21)
    B, s_txt = 2, 19
    shapes = [[(1, 8, 10), (1, 6, 14)]] * B
    s_img = 80 + 84
    lat = torch.randn(B, s_img, 64, generator=g)
    cond = torch.randn(B, s_txt, ocfg.joint_attention_dim, generator=g)
    mask = torch.ones(B, s_txt, dtype=torch.bool)
    t = torch.tensor([0.61, 0.61])
    ref = O.transformer_forward(sd, ocfg, lat, cond, mask, t.bfloat16().float(), shapes, [s_txt] * B)
    out = tr(hidden_states=lat, encoder_hidden_states=cond, encoder_hidden_states_mask=mask, timestep=t,
             img_shapes=shapes, txt_seq_lens=[s_txt] * B).sample
    assert O.rel_l2(out.float().cpu(), ref) < TOL
    assert O.rel_l2(out[:, 80:].float().cpu(), ref[:, 80:]) < TOL
    head = tr(hidden_states=lat, encoder_hidden_states=cond, encoder_hidden_states_mask=mask, timestep=t,
              img_shapes=shapes, txt_seq_lens=[s_txt] * B, _out_rows=80).sample
    assert head.shape == (B, 80, 64) and torch.equal(head, out[:, :80])
    with pytest.raises(ValueError):
        tr(hidden_states=lat, encoder_hidden_states=cond, timestep=t, img_shapes=[[(1, 8, 10)]] * B)


@pytest.mark.parametrize("graph,batched,cfg", [(False, True, True), (True, True, True), (False, False, True),
                                               (True, True, False)])
def test_edit_pipeline_per_step_latents(tiny, graph, batched, cfg):
    """pipe(image=<reference-image latents>, ...) against the oracle's edit loop, step by step."""
    import sys
    import os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    from make_golden import edit_inputs
    from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
    ocfg, sd, tr = tiny
    lat, ref_lat, cond, uncond, mask, shapes = edit_inputs(ocfg)
    neg = uncond if cfg else None
    ref, trace = O.edit_denoise_loop(sd, ocfg, lat, O.pack_latents(ref_lat), cond, mask, neg, mask if cfg else None,
                                     3, 4.0, img_shapes=shapes, return_all=True)
    pipe = B200QwenImagePipeline(tr)
    pipe.use_cuda_graph, pipe.batched_cfg = graph, batched
    out = pipe(image=ref_lat[:, :, 0], prompt_embeds=cond, prompt_embeds_mask=mask, negative_prompt_embeds=neg,
               negative_prompt_embeds_mask=mask if cfg else None, true_cfg_scale=4.0, num_inference_steps=3,
               height=256, width=256, latents=lat, output_type="latent", record_steps=True).images
    assert out.shape == ref.shape
    for i, (got, want) in enumerate(zip(pipe.last_step_latents, trace)):
        err = O.rel_l2(got.float().cpu(), want)
        assert err < TOL, f"step {i}: rel-L2 {err}"
    # the update itself, not just the latents: |final - initial| is 0.32 |initial| here, and the bf16 storage
    # of the latents alone is 0.004 |initial| = 1.3e-2 of the update (what the oracle run in bf16 shows too)
    assert O.rel_l2(out.float().cpu() - lat, ref - lat) < 5 * TOL
    if cfg and graph:
        gold = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tiny_widen.pt"))
        assert O.rel_l2(out.float().cpu(), gold["edit_final"]) < TOL


def test_reference_placeholder_loops(tiny):
    """The reference's own loops (runner_adapter.py:127-141, model.py:117-135) through the mirrored adapter and
    vLLM-facing wrapper, against their oracle restatement and the committed golden."""
    import os
    from vllm_omni_draft_b200.adapter import QwenImageCustomInputs, QwenImageRunnerAdapter
    from vllm_omni_draft_b200.gen_model import QwenImageGenModel, QwenImageInputs, QwenImageOutputMode
    ocfg, sd, tr = tiny
    gold = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tiny_widen.pt"))
    g = torch.Generator().manual_seed(5)
    lat64 = torch.randn(2, 64, 8, 12, generator=g)
    cond2 = torch.randn(2, 9, ocfg.joint_attention_dim, generator=g)
    mask2 = torch.ones(2, 9, dtype=torch.bool)
    ts = (torch.linspace(1.0, 0.0, steps=3) * 1000).long()
    model = QwenImageGenModel(transformer=tr, vae=None)
    # model.py:122-129 passes no img_shapes: a non-square grid cannot be derived -> ValueError
    with pytest.raises(ValueError):
        model.forward(qwen_inputs=QwenImageInputs(prompt_embeds=cond2, prompt_embeds_mask=mask2, image_latents=lat64,
                                                  timesteps=ts, output_mode=QwenImageOutputMode.LATENTS))
    sq = lat64[:, :, :, :8].contiguous()
    want = O.reference_gen_loop(sd, ocfg, sq, cond2, mask2, ts, step_size=0.1)
    got = model.forward(qwen_inputs=QwenImageInputs(prompt_embeds=cond2, prompt_embeds_mask=mask2, image_latents=sq,
                                                    timesteps=ts, guidance_scale=4.0,
                                                    output_mode=QwenImageOutputMode.LATENTS))
    assert got.shape == sq.shape and got.dtype == torch.bfloat16
    assert O.rel_l2(got.float().cpu(), want) < TOL
    # |update| is 0.1 |latents| (step size 0.1): bf16 storage of the latents is 4e-2 of it
    assert O.rel_l2(got.float().cpu() - sq, want - sq) < 12 * TOL
    with pytest.raises(RuntimeError):      # pixels requested, no VAE
        model.forward(qwen_inputs=QwenImageInputs(prompt_embeds=cond2, prompt_embeds_mask=mask2, image_latents=sq,
                                                  timesteps=ts))
    ad = QwenImageRunnerAdapter(transformer=tr, vae=None)
    inp = QwenImageCustomInputs(prompt_embeds=cond2, prompt_embeds_mask=mask2, image_latents=lat64,
                                img_shapes=[(1, 8, 12)], txt_seq_lens=[9, 9], num_inference_steps=3,
                                output_mode=QwenImageCustomInputs.QwenImageOutputMode.LATENTS)
    with pytest.raises(ValueError):        # the reference's own length check (types.py:103-106)
        ad.generate(inp, schedule="reference_placeholder")
    inp.img_shapes = [(1, 8, 12), (1, 8, 12)]      # one (F,H,W) per sample; upstream reads entry [0]
    got = ad.generate(inp, schedule="reference_placeholder")
    assert O.rel_l2(got.float().cpu(), gold["adapter_loop"]) < TOL
    assert O.rel_l2(got.float().cpu() - lat64, gold["adapter_loop"] - lat64) < 3 * TOL


def test_decoded_image_psnr(tiny):
    """north_star gate: the final decoded image reaches PSNR >= 40 dB against the reference path.  Both sides
    decode with the same (random-init, tiny) AutoencoderKLQwenImage-shaped decoder: the oracle's 5-D causal
    restatement in fp32 on the oracle's latents vs this package's one-frame bf16 decoder on the CUDA latents."""
    from oracle import vae_oracle as V
    from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
    from vllm_omni_draft_b200.vae import B200QwenImageVAEDecoder
    ocfg, sd, tr = tiny
    vcfg = V.tiny_vae_config()
    vsd = V.init_vae_state_dict(vcfg, seed=0)
    vae = B200QwenImageVAEDecoder(base_dim=vcfg.base_dim, num_res_blocks=vcfg.num_res_blocks, device="cuda:0")
    vae.load_diffusers_state_dict(vsd)
    lat, cond, uncond, mask, shapes = O.synthetic_inputs(ocfg, 256, 256, 32)
    ref_lat = O.denoise_loop(sd, ocfg, lat, cond, mask, uncond, mask, 4, 4.0, img_shapes=shapes)
    ref_img = V.decode_packed(vsd, vcfg, ref_lat, 32, 32)
    pipe = B200QwenImagePipeline(tr, vae=vae)
    img = pipe(prompt_embeds=cond, prompt_embeds_mask=mask, negative_prompt_embeds=uncond,
               negative_prompt_embeds_mask=mask, true_cfg_scale=4.0, num_inference_steps=4, height=256, width=256,
               latents=lat, output_type="pt").images
    assert img.shape == (1, 3, 256, 256) and img.dtype == torch.bfloat16
    want = (ref_img / 2 + 0.5).clamp(0, 1)
    db = V.psnr(img.float().cpu(), want, data_range=1.0)
    print(f"decoded-image PSNR vs oracle: {db:.1f} dB (image std {float(want.std()):.3f})")
    assert float(want.std()) > 0.02          # not a flat image: the comparison means something
    assert db >= 40.0
    # the decoder alone, same latents on both sides
    same = pipe.decode_latents(ref_lat.to("cuda:0", torch.bfloat16), 32, 32)
    assert V.psnr(same.float().cpu(), ref_img) >= 40.0


def test_vae_encoder_decoder_at_real_dimensions(cuda_device):
    """Rows b2 / f1 at the real AutoencoderKLQwenImage dimensions (base 96, mult (1,2,4,4), 2 residual blocks,
    z = 16; random-init): the one-frame bf16 encoder and decoder on the GPU against the 5-D causal fp32 restatement
    (oracle/vae_oracle.py) run on the same GPU, on a non-square 384 x 512 image."""
    from oracle import vae_oracle as V
    from vllm_omni_draft_b200.vae import B200QwenImageVAE
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    vcfg = V.VAEOracleConfig()
    vsd = {k: v.bfloat16().float().cuda() for k, v in V.init_vae_state_dict(vcfg, seed=1, weight_std=0.03).items()}
    vae = B200QwenImageVAE(device="cuda:0")
    vae.load_diffusers_state_dict(vsd)
    g = torch.Generator().manual_seed(12)
    x = (torch.rand(1, 3, 1, 384, 512, generator=g) * 2 - 1).bfloat16().float().cuda()
    want = V.encode(vsd, vcfg, x)
    dist = vae.encode(x).latent_dist
    got = torch.cat([dist.mean, dist.logvar], 1).float()
    assert got.shape == want.shape == (1, 32, 1, 48, 64)
    err_mean = O.rel_l2(dist.mean.float(), want[:, :16])
    err_all = O.rel_l2(got, want)
    z = want[:, :16].contiguous()
    img_want = V.decode(vsd, vcfg, z)
    img = vae.decode(z).sample.float()
    db = V.psnr(img, img_want)
    print(f"VAE real dims: encode mean rel-L2 {err_mean:.3e}, moments {err_all:.3e}; decode PSNR {db:.1f} dB "
          f"(image std {float(img_want.std()):.3f})")
    assert err_mean < 3e-2 and err_all < 3e-2          # bf16 activations through 11 residual blocks: 2.0e-2 measured
    assert float(img_want.std()) > 0.02 and db >= 40.0


def test_pipeline_takes_pixel_images_like_the_reference(tiny):
    """The reference's literal call (src/performance_enhancement/test_performance.py:244-251, 277):
    `pipe(image=<list of PIL images>, prompt_embeds=, prompt_embeds_mask=, generator=, num_inference_steps=,
    true_cfg_scale=1.0, num_images_per_prompt=1).images` -> PIL images at the edit pipeline's working size
    (~1024^2 at the input's aspect ratio), checked against the oracle: Lanczos resize, VAE encode (mode),
    normalise, pack, edit loop, decode."""
    import numpy as np
    from PIL import Image
    from oracle import vae_oracle as V
    from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
    from vllm_omni_draft_b200.vae import B200QwenImageVAE
    ocfg, sd, tr = tiny
    vcfg = V.tiny_vae_config()
    vsd = V.init_vae_state_dict(vcfg, seed=0)
    vae = B200QwenImageVAE(base_dim=vcfg.base_dim, num_res_blocks=vcfg.num_res_blocks, device="cuda:0")
    vae.load_diffusers_state_dict(vsd)
    pipe = B200QwenImagePipeline(tr, vae=vae)
    rng = np.random.default_rng(3)
    yy, xx = np.mgrid[0:480, 0:640]
    images = []
    for k in range(2):      # smooth synthetic photographs, 640 x 480
        base = np.stack([np.sin(xx / (40.0 + 10 * k) + c) * np.cos(yy / (55.0 - 5 * k) - c) for c in range(3)], -1)
        images.append(Image.fromarray(((base * 0.4 + 0.5 + rng.normal(0, 0.02, base.shape)).clip(0, 1) * 255).astype(np.uint8)))
    batch_size, mock_seq_len, steps = 2, 40, 3
    g = torch.Generator().manual_seed(4)
    prompt_embeds = torch.randn(batch_size, mock_seq_len, ocfg.joint_attention_dim, generator=g).to("cuda:0", torch.bfloat16)
    prompt_embeds_mask = torch.ones(batch_size, mock_seq_len, device="cuda:0", dtype=torch.bool)
    generator = torch.Generator(device="cuda:0").manual_seed(42)
    gen_kwargs = {"image": images, "prompt_embeds": prompt_embeds, "prompt_embeds_mask": prompt_embeds_mask,
                  "generator": generator, "num_inference_steps": steps, "true_cfg_scale": 1.0,
                  "num_images_per_prompt": 1}
    out = pipe(**gen_kwargs).images
    W, H = 1184, 896                                   # calculate_dimensions(1024*1024, 640/480)
    assert isinstance(out, list) and len(out) == 2 and all(im.size == (W, H) for im in out)
    # oracle side, fp32 on the host
    dev = "cpu"
    px = torch.stack([torch.from_numpy(np.asarray(im.resize((W, H), resample=Image.LANCZOS), dtype=np.float32) / 255.0)
                      for im in images]).permute(0, 3, 1, 2) * 2 - 1
    vsd_d = {k: v.to(dev) for k, v in vsd.items()}
    sd_d = {k: v.to(dev) for k, v in sd.items()}
    ref_lat = V.normalize_latents(V.encode_mode(vsd_d, vcfg, px.to(dev).unsqueeze(2)), vcfg)
    Hl, Wl = H // 8, W // 8
    noise = torch.randn(batch_size, 1, 16, Hl, Wl, generator=torch.Generator(device="cuda:0").manual_seed(42),
                        device="cuda:0", dtype=torch.bfloat16)      # the draw the call made (prompt dtype, generator's device)
    lat = O.pack_latents(noise.float().cpu())
    shapes = [[(1, Hl // 2, Wl // 2), (1, Hl // 2, Wl // 2)]] * batch_size
    final = O.edit_denoise_loop(sd_d, ocfg, lat, O.pack_latents(ref_lat), prompt_embeds.float().cpu(),
                                prompt_embeds_mask.cpu(), None, None, steps, 1.0, img_shapes=shapes)
    want = (V.decode_packed(vsd_d, vcfg, final, Hl, Wl) / 2 + 0.5).clamp(0, 1).cpu()
    got = torch.stack([torch.from_numpy(np.asarray(im, dtype=np.float32) / 255.0) for im in out]).permute(0, 3, 1, 2)
    db = V.psnr(got, want, data_range=1.0)
    print(f"pixel-image edit call vs oracle: PSNR {db:.1f} dB (image std {float(want.std()):.3f})")
    assert float(want.std()) > 0.02 and db >= 40.0


_MGPU_CASES = [
    # (gpus, env) — CFG-parallel at 2, CFG x Ulysses at 4/8, pure Ulysses (QI_CFG_PARALLEL=0), both exchanges,
    # ragged shards (272x240 px -> 255 image tokens, 33 prompt tokens) and batch 2
    (2, dict(QI_GRAPH="0")),
    (2, dict(QI_GRAPH="1")),
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="peer")),
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="nccl")),
    # arrival flags folded into the producer / consumer kernels instead of two barrier launches per block
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="peer", QI_PEER_FOLD="1")),
    # custom `sigmas=` through the launcher
    (2, dict(QI_GRAPH="1", QI_PAR_SIGMAS="1")),
    # scatter through 16-byte st.global instead of TMA bulk tensor stores
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="peer", QI_SCATTER_TMA="0")),
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="peer", QI_PAR_H="272", QI_PAR_W="240",
             QI_PAR_TXT="33", QI_PAR_BATCH="2")),
    # edit mode (reference-image tokens sharded beside the noise tokens): pure Ulysses and CFG-parallel
    (2, dict(QI_GRAPH="1", QI_CFG_PARALLEL="0", QI_SP_EXCHANGE="peer", QI_PAR_EDIT="1", QI_PAR_TXT="33")),
    (2, dict(QI_GRAPH="1", QI_PAR_EDIT="1")),
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


def test_pipeline_sessions_follow_weight_reloads_and_cache_limit(cuda_device):
    """(1) a cached session (hoisted modulation table + captured graph) must not survive a weight
    reload or a flip of the kernel switches it was captured under — the same-shape call after the reload has to
    match the oracle on the NEW weights; (2) max_cached_sessions = 0 means "keep nothing", not a crash."""
    from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
    ocfg = O.tiny_config()
    sd_a = O.init_state_dict(ocfg, seed=0, bias_std=0.02, norm_weight_std=0.1)
    sd_b = O.init_state_dict(ocfg, seed=5, bias_std=0.02, norm_weight_std=0.1)
    tr = _build(ocfg, sd_a)
    lat, cond, uncond, mask, shapes = O.synthetic_inputs(ocfg, 256, 256, 32)
    kw = dict(prompt_embeds=cond, prompt_embeds_mask=mask, negative_prompt_embeds=uncond,
              negative_prompt_embeds_mask=mask, true_cfg_scale=4.0, num_inference_steps=3, height=256, width=256,
              latents=lat, output_type="latent")
    ref_a = O.denoise_loop(sd_a, ocfg, lat, cond, mask, uncond, mask, 3, 4.0, img_shapes=shapes)
    ref_b = O.denoise_loop(sd_b, ocfg, lat, cond, mask, uncond, mask, 3, 4.0, img_shapes=shapes)
    assert O.rel_l2(ref_a, ref_b) > 0.05            # the two weight sets are tell-apart-able
    pipe = B200QwenImagePipeline(tr)
    pipe.use_cuda_graph = True
    assert O.rel_l2(pipe(**kw).images.float().cpu(), ref_a) < TOL
    first = next(iter(pipe._sessions.values()))
    assert O.rel_l2(pipe(**kw).images.float().cpu(), ref_a) < TOL and next(iter(pipe._sessions.values())) is first
    tr.load_diffusers_state_dict(sd_b)              # same shapes, new weights
    assert O.rel_l2(pipe(**kw).images.float().cpu(), ref_b) < TOL
    assert len(pipe._sessions) == 1 and next(iter(pipe._sessions.values())) is not first
    second = next(iter(pipe._sessions.values()))
    tr.fuse_qknorm_rope = not tr.fuse_qknorm_rope   # a different kernel sequence than the captured one
    assert O.rel_l2(pipe(**kw).images.float().cpu(), ref_b) < TOL
    assert next(iter(pipe._sessions.values())) is not second
    tr.fuse_qknorm_rope = not tr.fuse_qknorm_rope
    pipe.clear_sessions()
    pipe.max_cached_sessions = 0
    for _ in range(2):
        assert O.rel_l2(pipe(**kw).images.float().cpu(), ref_b) < TOL
        assert not pipe._sessions
