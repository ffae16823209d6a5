"""Opt-in compute-sanitizer pass: racecheck + synccheck over the tiny config's kernels — the code has
exactly what those tools exist for (mbarrier phase tracking, cross-CTA multicast commits, TMEM hand-offs between
warp roles).  Slow (every kernel runs instrumented), so it only runs with QI_SANITIZER=1:

    QI_SANITIZER=1 python -m pytest tests/test_sanitizer_gpu.py -m gpu -s        (tools/gpu_run.sh sanitizer)
"""
import os
import shutil
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.sanitizer]

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_SCRIPT = r"""
import sys, torch
sys.path.insert(0, %r)
from oracle import qwen_image_oracle as O
from vllm_omni_draft_b200.config import MMDiTConfig
from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
from vllm_omni_draft_b200.transformer import B200QwenImageTransformer2DModel
import os
os.environ["QI_GEMM_TUNE"] = "0"
ocfg = O.tiny_config()
sd = O.init_state_dict(ocfg, seed=0, bias_std=0.02, norm_weight_std=0.1)
lat, cond, uncond, mask, shapes = O.synthetic_inputs(ocfg, 128, 128, 32)
tr = B200QwenImageTransformer2DModel(MMDiTConfig.tiny(), device="cuda:0")
tr.load_diffusers_state_dict(sd)
out = B200QwenImagePipeline(tr)(prompt_embeds=cond, prompt_embeds_mask=mask, negative_prompt_embeds=uncond,
    negative_prompt_embeds_mask=mask, true_cfg_scale=4.0, num_inference_steps=1, height=128, width=128, latents=lat,
    output_type="latent").images
torch.cuda.synchronize()
ref = O.denoise_loop(sd, ocfg, lat, cond, mask, uncond, mask, 1, 4.0, img_shapes=shapes)
print("REL_L2", O.rel_l2(out.float().cpu(), ref))
""" % ROOT


@pytest.mark.parametrize("tool", ["racecheck", "synccheck"])
def test_tiny_step_under_compute_sanitizer(cuda_device, tool):
    if os.environ.get("QI_SANITIZER") != "1":
        pytest.skip("opt-in: set QI_SANITIZER=1")
    cs = shutil.which("compute-sanitizer") or "/usr/local/cuda/bin/compute-sanitizer"
    if not os.path.exists(cs):
        pytest.skip("compute-sanitizer not installed")
    p = subprocess.run([cs, "--tool", tool, "--error-exitcode", "77", "--kernel-regex", "kns=qi", sys.executable, "-c", _SCRIPT],
                       capture_output=True, text=True, timeout=1400)
    tail = (p.stdout + p.stderr)[-3000:]
    print(tail)
    if "compute-sanitizer is closed on this pool" in tail:
        pytest.skip("compute-sanitizer is disabled by the GPU pool's wrapper (not a finding about the kernels)")
    assert p.returncode == 0, f"{tool} reported errors:\n{tail}"
    assert "REL_L2" in p.stdout
