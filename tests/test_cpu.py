"""CPU suite (-m "not gpu"): oracle vs committed golden vectors, bit-exact integer tables through the
C ABI, host logic (scheduler, packing, validation), and the shared library's exported symbols."""
import ctypes
import hashlib
import os
import re
import sys

import pytest
import torch

from oracle import qwen_image_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def sha(t):
    return hashlib.sha256(t.contiguous().numpy().tobytes()).hexdigest()


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    from vllm_omni_draft_b200 import _lib
    return _lib


# ------------------------------------------------------------------ oracle vs golden
def test_oracle_reproduces_golden_forward_and_loop():
    gold = torch.load(os.path.join(GOLD, "tiny_cfg1.pt"))
    cfg = O.tiny_config()
    sd = O.init_state_dict(cfg, seed=0, bias_std=0.02, norm_weight_std=0.1)
    lat, cond, uncond, mask, shapes = O.synthetic_inputs(cfg, 256, 256, 32)
    fwd = O.transformer_forward(sd, cfg, lat, cond, mask, torch.tensor([0.7446]).bfloat16().float(), shapes, [32])
    assert O.rel_l2(fwd, gold["forward_sample"]) < 1e-5
    final, trace = O.denoise_loop(sd, cfg, lat, cond, mask, uncond, mask, 4, 4.0, img_shapes=shapes, return_all=True)
    assert O.rel_l2(torch.stack(trace), gold["step_latents"]) < 1e-5
    assert O.rel_l2(final, gold["final"]) < 1e-5


def test_oracle_scheduler_known_answers():
    gold = torch.load(os.path.join(GOLD, "tiny_cfg1.pt"))
    ts, sig = O.flowmatch_sigmas(4, 256)
    assert torch.equal(sig, gold["sigmas4"]) and torch.equal(ts, gold["timesteps4"])
    # closed-form facts of the Qwen-Image schedule: starts at 1, terminal sigma 0.02, trailing 0
    ts50, sig50 = O.flowmatch_sigmas(50, 4096)
    assert sig50[0] == 1.0 and abs(float(sig50[-2]) - 0.02) < 1e-6 and sig50[-1] == 0
    assert torch.all(sig50[1:] < sig50[:-1])
    assert abs(O.calculate_shift(4096) - 0.693548) < 1e-5      # mu ~= 0.6935 at 4096 tokens
    assert torch.equal(sig50, gold["sigmas50_4096"])


def test_oracle_bf16_tracks_fp32():
    cfg = O.tiny_config()
    sd = O.init_state_dict(cfg, seed=0, bias_std=0.02, norm_weight_std=0.1)
    lat, cond, uncond, mask, shapes = O.synthetic_inputs(cfg, 256, 256, 32)
    ref = O.denoise_loop(sd, cfg, lat, cond, mask, uncond, mask, 2, 4.0, img_shapes=shapes)
    sdb = {k: v.bfloat16() for k, v in sd.items()}
    got = O.denoise_loop(sdb, cfg, lat.bfloat16(), cond.bfloat16(), mask, uncond.bfloat16(), mask, 2, 4.0,
                         img_shapes=shapes)
    assert O.rel_l2(got.float(), ref) < 1e-2


def test_oracle_rope_and_pack_properties():
    # rotation preserves the norm of every (even, odd) pair; pack/unpack is a bijection
    cfg = O.tiny_config()
    img_f, txt_f = O.rope_freqs(cfg, [(1, 5, 7)], 9)
    x = torch.randn(1, 35, 4, 64)
    y = O.apply_rope(x, img_f)
    assert torch.allclose(x.view(1, 35, 4, 32, 2).norm(dim=-1), y.view(1, 35, 4, 32, 2).norm(dim=-1), atol=1e-5)
    lat = torch.randn(2, 16, 1, 10, 14)
    assert torch.equal(O.unpack_latents(O.pack_latents(lat.transpose(1, 2)), 10, 14), lat)
    with pytest.raises(ValueError):
        O.rope_position_ids([(1, 128, 128)], 4096)


# ------------------------------------------------------------------ C ABI: load + symbols
def test_library_exports_every_declared_symbol(built):
    header = open(os.path.join(ROOT, "include", "qi_b200.h")).read()
    declared = set(re.findall(r"\b(qi_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    lib = ctypes.CDLL(built.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in qi_b200.h but not exported"
    assert declared == set(built.EXPORTED_SYMBOLS)
    assert built.lib().qi_version() == 101


def test_attention_variant_switch_validates_its_argument(built):
    """qi_attention_set_variant is host-only state: the four shipped softmax forms (include/qi_b200.h: -1 = default,
    EMU in {0, 3} + 8 * SPLIT) are accepted, anything else — including the experiment variants that now live in
    tools/experiments — is QI_ERR_ARG, and the default is restored afterwards."""
    lib = built.lib()
    try:
        for v in (-1, 0, 3, 8, 11):
            assert lib.qi_attention_set_variant(v) == 0
        for v in (-2, 2, 9, 19, 43, 59, 67, 84, 85, 1000):
            assert lib.qi_attention_set_variant(v) != 0
    finally:
        assert lib.qi_attention_set_variant(-1) == 0
    assert not hasattr(lib, "qi_attention_set_debug")   # measurement hooks are not in the product library


def test_compute_entry_points_fail_loudly_without_gpu(built):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from vllm_omni_draft_b200 import _lib
    p = _lib.GemmProblem()
    assert built.lib().qi_gemm_bf16(ctypes.byref(p), 1, None) != 0
    assert built.lib().qi_device_sm_count() < 0
    with pytest.raises(RuntimeError):
        _lib.check(-2, "qi_gemm_bf16")


# ------------------------------------------------------------------ bit-exact integer tables
def test_rope_position_ids_bit_exact(built):
    from vllm_omni_draft_b200 import ops
    tables = torch.load(os.path.join(GOLD, "int_tables.pt"))
    for (f, h, w), txt in (((1, 16, 16), 32), ((1, 64, 64), 512), ((1, 83, 83), 512), ((1, 128, 128), 512)):
        ip, tp = ops.rope_position_ids([(f, h, w)], txt)
        oi, ot = O.rope_position_ids([(f, h, w)], txt)
        assert torch.equal(ip.long(), oi) and torch.equal(tp.long(), ot)
        assert sha(torch.cat([tp, ip])) == tables[f"rope_sha_{h}x{w}_{txt}"]
    ip, tp = ops.rope_position_ids([(1, 16, 16)], 32)
    assert torch.equal(torch.cat([tp, ip]), tables["rope_pos_16x16_32"])
    ip, tp = ops.rope_position_ids([(1, 8, 8), (1, 6, 10)], 7)
    assert torch.equal(torch.cat([tp, ip]), tables["rope_pos_two_images"])
    with pytest.raises(ValueError):
        ops.rope_position_ids([(1, 128, 128)], 4096)      # 64 + 4096 > 4096-entry table
    ops.rope_position_ids([(1, 128, 128)], 4032)          # the reference's "max 4064"-style bound


def test_pack_index_bit_exact(built):
    from vllm_omni_draft_b200 import ops
    from vllm_omni_draft_b200.pipeline import pack_latents, unpack_latents
    tables = torch.load(os.path.join(GOLD, "int_tables.pt"))
    for hl in (32, 128, 166):
        idx = ops.pack_latents_index(16, hl, hl)
        assert sha(idx) == tables[f"pack_sha_{hl}"]
        src = torch.arange(16 * hl * hl, dtype=torch.int64).view(1, 16, hl, hl)
        assert torch.equal(pack_latents(src).reshape(-1), idx)
    assert torch.equal(ops.pack_latents_index(16, 32, 32).int(), tables["pack_idx_32"])
    x = torch.randn(3, 16, 1, 12, 20)
    assert torch.equal(unpack_latents(pack_latents(x.transpose(1, 2)), 12, 20), x)
    with pytest.raises(ValueError):
        ops.pack_latents_index(16, 31, 32)


def test_joint_rope_table_matches_oracle_bitwise(built):
    from vllm_omni_draft_b200 import rope
    for cfg, shapes, txt in ((O.tiny_config(), [(1, 16, 16)], 32), (O.qwen_image_20b_config(), [(1, 9, 13)], 21)):
        tab = rope.joint_cos_sin(cfg.axes_dims_rope, cfg.rope_theta, shapes, txt)
        i, t = O.rope_freqs(cfg, shapes, txt)
        assert torch.equal(tab, torch.view_as_real(torch.cat([t, i], 0)))
    assert rope.normalize_img_shapes([[(1, 4, 4)], [(1, 4, 4)]]) == ((1, 4, 4),)
    assert rope.normalize_img_shapes([(1, 4, 4)]) == ((1, 4, 4),)


# ------------------------------------------------------------------ host logic
def test_schedule_matches_oracle():
    from vllm_omni_draft_b200.pipeline import FlowMatchEulerSchedule, calculate_shift, latent_hw
    for n, s in ((4, 256), (50, 4096), (50, 6889), (28, 16384), (2, 4096)):
        sch = FlowMatchEulerSchedule(n, s)
        ts, sig = O.flowmatch_sigmas(n, s)
        assert torch.equal(sch.sigmas, sig) and torch.equal(sch.timesteps, ts)
        assert torch.equal(sch.dts, sig[1:] - sig[:-1])
    assert calculate_shift(4096) == O.calculate_shift(4096)
    assert latent_hw(1024, 1024) == (128, 128) and latent_hw(1328, 1328) == (166, 166)
    with pytest.raises(ValueError):
        FlowMatchEulerSchedule(0, 256)
    # caller-supplied pre-shift sigmas (upstream's `sigmas=` argument of the pipeline call)
    custom = [1.0, 0.8, 0.55, 0.3, 0.1]
    sch = FlowMatchEulerSchedule(5, 1024, sigmas=custom)
    ts, sig = O.flowmatch_sigmas(5, 1024, sigmas=custom)
    assert torch.equal(sch.sigmas, sig) and torch.equal(sch.timesteps, ts) and sig[0] == 1.0
    with pytest.raises(ValueError):
        FlowMatchEulerSchedule(4, 1024, sigmas=custom)


def test_config_and_input_validation():
    from vllm_omni_draft_b200.config import MMDiTConfig, RunConfig
    from vllm_omni_draft_b200.adapter import QwenImageCustomInputs, validate_custom_inputs
    assert MMDiTConfig().inner_dim == 3072 and MMDiTConfig.tiny().inner_dim == 256
    with pytest.raises(ValueError):
        RunConfig.from_env(bogus=1)
    ok = QwenImageCustomInputs(prompt_embeds=torch.zeros(1, 4, 8), prompt_embeds_mask=torch.ones(1, 4),
                               image_latents=torch.zeros(1, 16, 4, 4))
    validate_custom_inputs(ok)
    for bad in (dict(prompt_embeds=torch.zeros(4, 8)), dict(prompt_embeds_mask=torch.ones(4)),
                dict(image_latents=torch.zeros(16, 4, 4)), dict(prompt_embeds_mask=torch.ones(2, 4))):
        kw = dict(prompt_embeds=torch.zeros(1, 4, 8), prompt_embeds_mask=torch.ones(1, 4),
                  image_latents=torch.zeros(1, 16, 4, 4))
        kw.update(bad)
        with pytest.raises(ValueError):
            validate_custom_inputs(QwenImageCustomInputs(**kw))
    with pytest.raises(ValueError):
        validate_custom_inputs(QwenImageCustomInputs(prompt_embeds=torch.zeros(1, 4, 8),
                                                     prompt_embeds_mask=torch.ones(1, 4),
                                                     image_latents=torch.zeros(1, 16, 4, 4),
                                                     img_shapes=[(1, 2, 2)], txt_seq_lens=[4, 4]))


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the reference's CPU path = the oracle on the host cores) prints ONE JSON line
    with the keys the driver reads; runs without a GPU and without the CUDA library."""
    import json
    import subprocess
    import sys
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == "denoise steps/s" and j["unit"] == "steps/s"
    assert j["higher_is_better"] is True and j["vs_baseline"] is None and j["n_gpus"] == 1 and j["steps"] == 1
    assert j["value"] > 0 and abs(j["value"] - 1e3 / j["ms_per_step"]) < 1e-9
    assert j["cpu_baseline"]["kind"] == "port" and j["cpu_baseline"]["cores"] >= 1 and j["cpu_baseline"]["sample"]
    assert j["cpu_baseline"]["value"] == j["value"]
    assert j["e2e"] == {"value": j["value"], "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "4096 image tokens" in j["config"]["workload"] and "model" not in j["config"]
    # under torchrun only rank 0 prints; the other ranks exit 0 without work
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, env=dict(os.environ, RANK="1", WORLD_SIZE="2"))
    assert r.returncode == 0 and not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]


def test_static_barrier_size_limit_is_checked_before_any_cuda_call(built):
    """qi_peer_barrier_static (the entry point for hosts that cannot allocate device memory) keeps its pointer table
    in a fixed array; a group larger than that table is refused up front, with a message, GPU or not."""
    lib = built.lib()
    table = (ctypes.c_void_p * 9)(*([1] * 9))
    assert lib.qi_peer_barrier_static(table, 9, 0, 1, None, 2000, None) == -1
    assert b"n_peers = 9" in lib.qi_last_cuda_error()


def test_shipped_schedules_are_pinned_to_this_gemm_source(built):
    """schedules timed on an older build must not be reused silently.  The in-tree file carries the
    hash of the GEMM sources it was measured on and only keys of the current qi_version(); changing qi_gemm.cu /
    qi_common.cuh without bumping the version, re-tuning and re-pinning (tools/pin_schedules.py) fails here."""
    import json
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import pin_schedules
    with open(os.path.join(ROOT, "vllm_omni_draft_b200", "tuned_schedules.json")) as fh:
        shipped = json.load(fh)
    assert shipped.pop(pin_schedules.PIN_KEY) == pin_schedules.gemm_source_sha256(), \
        "GEMM source changed: bump qi_version(), re-tune on a B200, run tools/pin_schedules.py"
    tag = f"/v{built.lib().qi_version()}/"
    assert shipped and all(tag in k for k in shipped)
    from vllm_omni_draft_b200 import ops
    assert not any(k.startswith("__") for k in ops.tuned_schedules())
