#!/usr/bin/env python
"""bench.py — denoise steps/s of the Qwen-Image hot path (BASELINE.json metric) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path (oracle)

One "step" = everything between two successive latents with true CFG: cond forward + uncond forward
of the 60-block MMDiT + the fused CFG/Euler kernel.  Workload at N=1: BASELINE.json configs[1]
(Qwen-Image 20B, 1024x1024 -> 4096 image tokens, 512 text tokens, bf16, random-init, synthetic
prompt embeddings and latents).  N=2: CFG-parallel (cond / uncond ranks); N=4/8: CFG-parallel x
Ulysses sequence parallel (strong scaling of the same job).  Prints ONE JSON line on rank 0.

What is measured, in the order it runs:
  1. `e2e` (the headline): `B200QwenImagePipeline.__call__(prompt_embeds=<pinned host>, negative_prompt_embeds=...,
     latents=<pinned host>, num_inference_steps=50, output_type="latent")` returning to a pinned host tensor — the
     call the reference makes (test_performance.py:277), twice: the first call pays session setup (modulation table,
     GEMM autotuning, CUDA-graph capture; reported as `setup_ms`), the second is the number (`s_per_image`).
  2. `value`: K scheduler steps with everything resident in HBM (CUDA events, max over ranks), clocks sampled.
  3. `parity`: two blocks of the same width, same shapes, same parallel layout (same peer exchange, same graph
     path), 2 true-CFG steps, against the fp32 oracle on rank 0's GPU — the correctness of the layout being timed.
  4. `roofline` / `kernels`: every hot kernel alone (CUDA graph of 24 launches) with library bars beside it.
  5. `cpu_baseline`: the oracle on the host cores.
"""
from __future__ import annotations

import argparse
import csv
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

D, L, HEADS, DH, D_TXT = 3072, 60, 24, 128, 3584
TRAFFIC_CSV = os.path.join(ROOT, "profiles", "r02_ncu_mlp_up_batch2.csv")

WORKLOADS = {
    # name: (height, width, s_txt, batch) — BASELINE.json configs[1], [3], [4] and the tiny config[0]
    "cfg2": (1024, 1024, 512, 1),
    "cfg4": (2048, 2048, 512, 1),
    "cfg5": (1328, 1328, 512, 8),
    "tiny": (256, 256, 32, 1),
}


def algorithmic_flops_per_forward(s_img: int, s_txt: int, batch: int = 1, layers: int = L, d: int = D,
                                  d_txt: int = D_TXT) -> float:
    """per block 24 d^2 per token of GEMM + 4 S^2 d of attention (+ in/out projections)."""
    S = s_img + s_txt
    per_block = 24.0 * d * d * S * batch + 4.0 * S * S * d * batch + 24.0 * d * d * batch
    return layers * per_block + 2.0 * batch * (64 * d * s_img + d_txt * d * s_txt + 64 * d * s_img)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            j = json.load(f)
        return j.get("bf16_tflops", 1590.0), j.get("bf16_tflops_sustained", 1400.0), j.get("hbm_gbs", 6650.0), "measured"
    return 1590.0, 1400.0, 6650.0, "fallback"


def committed_traffic():
    """DRAM bytes per launch of the dominant kernel, read from the committed `ncu --set full` capture of that launch
    (tools/ncu_summary.py wrote the CSV).  Fails loudly if the capture is missing or is of another kernel."""
    if not os.path.exists(TRAFFIC_CSV):
        raise RuntimeError(f"{TRAFFIC_CSV} is missing: capture it with tools/gpu_run.sh ncu_mlp_up (ncu --set full)")
    rd = wr = None
    name = None
    with open(TRAFFIC_CSV) as f:
        for row in csv.DictReader(l for l in f if not l.startswith("#")):
            if "gemm_pair_bf16_kernel" not in row.get("kernel", "") or row["kernel"].rsplit("#", 1)[-1].isdigit():
                continue        # another kernel, or a per-launch row (the averaged row is the one used)
            name = row["kernel"]
            if row["metric"] == "dram__bytes_read.sum":
                rd = float(row["value"])
            elif row["metric"] == "dram__bytes_write.sum":
                wr = float(row["value"])
    if rd is None or wr is None:
        raise RuntimeError(f"{TRAFFIC_CSV} holds no dram__bytes_read/write rows for gemm_pair_bf16_kernel")
    return rd + wr, name


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region."""

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_reference_sample(s_img: int, s_txt: int, n_blocks: int = 2, repeats: int = 1):
    """The reference has no CPU-specific code: its CPU path is the same PyTorch graph on
    device='cpu' (test_performance.py:25).  Bounded sample of the cfg-2 workload: `n_blocks` of the
    60 MMDiT blocks of ONE conditional forward at full width/sequence, fp32, all host threads;
    extrapolated to a step as x (60 / n_blocks) x 2 forwards."""
    from oracle import qwen_image_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = O.OracleConfig(num_layers=n_blocks)
    sd = O.init_state_dict(cfg, seed=0)
    g = torch.Generator().manual_seed(42)
    h = torch.randn(1, s_img, D, generator=g)
    e = torch.randn(1, s_txt, D, generator=g)
    temb = torch.randn(1, D, generator=g)
    side = int(s_img ** 0.5)
    img_f, txt_f = O.rope_freqs(cfg, [(1, side, s_img // side)], s_txt)
    times = []
    with torch.no_grad():
        for _ in range(repeats):
            t0 = time.perf_counter()
            hh, ee = h, e
            for i in range(n_blocks):
                ee, hh = O.block_forward(sd, i, cfg, hh, ee, temb, img_f, txt_f)
            times.append(time.perf_counter() - t0)
    return times, torch.get_num_threads()


def cpu_cfg1_loop(timed: int = 5):
    """SURVEY §8d / BASELINE.md §4: the oracle at BASELINE.json configs[0] (tiny MMDiT, 256x256, 4 FlowMatch steps,
    batch 1, fp32, true CFG) on all host threads: 1 warm-up + `timed` full 4-step loops, median.  Not extrapolated."""
    from oracle import qwen_image_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = O.tiny_config()
    sd = O.init_state_dict(cfg, seed=0)
    lat, cond, uncond, mask, shapes = O.synthetic_inputs(cfg, 256, 256, 32)
    ts = []
    with torch.no_grad():
        for i in range(timed + 1):
            t0 = time.perf_counter()
            O.denoise_loop(sd, cfg, lat, cond, mask, uncond, mask, num_inference_steps=4, true_cfg_scale=4.0,
                           img_shapes=shapes)
            if i:
                ts.append(time.perf_counter() - t0)
    med = statistics.median(ts)
    return {"workload": "BASELINE.json configs[0]: tiny MMDiT (2 blocks, d=256, 4 heads), 256x256, 4 true-CFG FlowMatch steps, fp32",
            "steps_per_s": 4.0 / med, "s_per_image": med, "loops_timed": timed, "statistic": "median", "extrapolated": False}


class _AliasedBlocks(dict):
    """State dict of `distinct` MMDiT blocks that answers for all 60: block i reads block i % distinct.  A block's
    fp32 weights are 1.36 GB, far beyond any host cache, so reuse does not make the CPU run faster than 60 distinct
    blocks would; it keeps the fp32 model at a few GB of host memory instead of 82 GB."""

    def __init__(self, base, distinct):
        super().__init__(base)
        self.distinct = distinct

    def __getitem__(self, key):
        if key.startswith("transformer_blocks."):
            head, idx, rest = key.split(".", 2)
            key = f"{head}.{int(idx) % self.distinct}.{rest}"
        return super().__getitem__(key)


def cpu_full_step(s_img: int, s_txt: int, distinct: int = 4) -> dict:
    """ONE whole true-CFG denoising step of the cfg-2 workload on the host cores, nothing extrapolated: both
    forwards through all 60 block evaluations with img_in / txt_norm / txt_in / time embedding / norm_out /
    proj_out, the CFG combine with norm rescale and the FlowMatch Euler update (fp32 oracle, all host threads).
    What the extrapolated `cpu_baseline.value` leaves out is inside this number."""
    from oracle import qwen_image_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = O.OracleConfig(num_layers=L)
    sd = _AliasedBlocks(O.init_state_dict(O.OracleConfig(num_layers=distinct), seed=0), distinct)
    g = torch.Generator().manual_seed(42)
    side = int(s_img ** 0.5)
    lat = torch.randn(1, s_img, 64, generator=g)
    cond = torch.randn(1, s_txt, 3584, generator=g)
    uncond = torch.randn(1, s_txt, 3584, generator=g)
    mask = torch.ones(1, s_txt, dtype=torch.bool)
    shapes = [[(1, side, s_img // side)]]
    timesteps, sigmas = O.flowmatch_sigmas(50, s_img)
    with torch.no_grad():
        t0 = time.perf_counter()
        t = (timesteps[0] / 1000).expand(1)
        c = O.transformer_forward(sd, cfg, lat, cond, mask, t, shapes, [s_txt])
        u = O.transformer_forward(sd, cfg, lat, uncond, mask, t, shapes, [s_txt])
        out = O.scheduler_step(O.true_cfg_combine(c, u, 4.0), lat, sigmas[0], sigmas[1])
        dt = time.perf_counter() - t0
    assert torch.isfinite(out).all()
    return {"full_step_s": dt, "steps_per_s": 1.0 / dt, "cores": torch.get_num_threads(), "extrapolated": False,
            "what": f"one true-CFG step, 2 forwards x {L} block evaluations + in/out projections, time embedding, "
                    f"CFG combine, Euler update; fp32 oracle; {distinct} distinct blocks' weights reused (host memory)"}


def cpu_baseline_block(s_img, s_txt, batch, steps_per_image, repeats, warm):
    """cfg-2 (or the named workload) on the host cores: bounded sample, median, extrapolated and labelled so."""
    n_blocks = 2
    times, threads = cpu_reference_sample(s_img, s_txt, n_blocks, repeats=warm + repeats)
    t_sample = statistics.median(times[warm:])
    step_s = t_sample * (L / n_blocks) * 2 * batch
    return {"value": 1.0 / step_s, "unit": "steps/s", "cores": threads, "kind": "port", "extrapolated": True,
            "statistic": f"median of {repeats} after {warm} warm-up",
            "sample": f"{n_blocks} of {L} MMDiT blocks of one cond forward at the workload's full width and sequence "
                      f"(batch 1), fp32 oracle on all host threads, extrapolated x{L // n_blocks} blocks x2 forwards"
                      + (f" x{batch} batch" if batch > 1 else ""),
            "s_per_image": step_s * steps_per_image,
            "extrapolation_check": "profiles/r03_cpu_full_step.json (bench.py --impl reference --full-step, 16 cores): ONE "
                                   "whole cfg-2 step measured 75.56 s vs 75.74 s extrapolated from 2 blocks (0.2 % apart)"}, step_s


def run_reference_arm(args, workload):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, step_s = cpu_baseline_block(workload["s_img"], workload["s_txt"], workload["batch"],
                                    workload["num_inference_steps"], repeats=max(args.steps, 1), warm=max(args.warmup, 1))
    cb["cfg1_loop"] = cpu_cfg1_loop()
    if args.full_step and workload["batch"] == 1:
        fs = cpu_full_step(workload["s_img"], workload["s_txt"])
        fs["extrapolated_over_measured"] = fs["full_step_s"] / step_s
        cb["full_step"] = fs
    value = cb["value"]
    line = {"impl": "reference", "metric": "denoise steps/s", "value": value, "unit": "steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload["config"], "cpu_baseline": cb,
            "e2e": {"value": value, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "s_per_image": cb["s_per_image"], "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ GPU arm
def make_workload(name: str, gpus: int):
    H, W, s_txt, B = WORKLOADS[name]
    s_img = (H // 16) * (W // 16)
    what = (f"Qwen-Image 20B MMDiT (60 blocks, d=3072, 24x128 heads) random-init, {H}x{W} ({s_img} image tokens), "
            f"{s_txt} text tokens, batch {B}, true CFG (2 forwards/step), 50-step FlowMatch schedule"
            if name != "tiny" else "tiny MMDiT (2 blocks, d=256) 256x256, 32 text tokens, true CFG, 50-step schedule")
    return {"name": name, "H": H, "W": W, "s_img": s_img, "s_txt": s_txt, "batch": B, "num_inference_steps": 50,
            "config": {"workload": what,
                       "parallelism": {1: "single", 2: "cfg2", 4: "cfg2 x ulysses2", 8: "cfg2 x ulysses4"}.get(gpus, "?"),
                       "l2": "weights of 60 distinct blocks (40.9 GB) stream through every step: inputs larger than L2, no flush needed",
                       # storage of the GPU arm's residual streams (set here so both arms print the same config)
                       "residual_streams": "bfloat16" if os.environ.get("QI_RESIDUAL", "fp32") == "bf16" else "float32"}}


def layout_parity(world, rank, dev, wl, use_graph, tiny):
    """Correctness of the layout being timed: a 2-block model of the SAME width run through the SAME host path
    (pipeline call -> ParallelDenoiser -> peer exchange kernels -> CUDA graph) at the SAME shapes for 4 true-CFG
    steps, against the oracle's loop in fp32 on rank 0's GPU (weights read from rank 0's replica); with more than
    one GPU also against the same call on rank 0 alone (what the sharding itself changes)."""
    from oracle import qwen_image_oracle as O
    from vllm_omni_draft_b200.config import MMDiTConfig
    from vllm_omni_draft_b200.launcher import ParallelDenoiser
    steps = 4
    cfg = MMDiTConfig.tiny() if tiny else MMDiTConfig(num_layers=2)
    ocfg = O.tiny_config() if tiny else O.OracleConfig(num_layers=2)
    den = ParallelDenoiser(cfg, dev, world, rank, seed=7, use_cuda_graph=use_graph)
    den.tr.init_random_(7, bias_std=0.02, norm_weight_std=0.1)      # same seed on every rank: replicated weights
    den.pipe.use_cuda_graph = use_graph
    B, s_img, s_txt = wl["batch"], wl["s_img"], wl["s_txt"]
    g = torch.Generator().manual_seed(123)
    lat = torch.randn(B, s_img, 64, generator=g).bfloat16()
    cond = torch.randn(B, s_txt, cfg.joint_attention_dim, generator=g).bfloat16()
    uncond = torch.randn(B, s_txt, cfg.joint_attention_dim, generator=g).bfloat16()
    mask = torch.ones(B, s_txt, dtype=torch.bool)
    out = den.pipe(prompt_embeds=cond, prompt_embeds_mask=mask, negative_prompt_embeds=uncond,
                   negative_prompt_embeds_mask=mask, true_cfg_scale=4.0, num_inference_steps=steps, latents=lat,
                   height=wl["H"], width=wl["W"], output_type="latent").images
    torch.cuda.synchronize()
    res = None
    if rank == 0:
        with torch.no_grad():
            hh, ww = wl["H"] // 16, wl["W"] // 16
            shapes = [[(1, hh, ww)]] * B
            sd = O.ParamView(den.tr)
            timesteps, sigmas = O.flowmatch_sigmas(steps, s_img)
            x, c_d, u_d, m_d = lat.to(dev).float(), cond.to(dev).float(), uncond.to(dev).float(), mask.to(dev)
            for i, t in enumerate(timesteps):
                t_in = (t.to(torch.bfloat16) / 1000).float().expand(B).to(dev)
                c = O.transformer_forward(sd, ocfg, x, c_d, m_d, t_in, shapes, [s_txt] * B)
                u = O.transformer_forward(sd, ocfg, x, u_d, m_d, t_in, shapes, [s_txt] * B)
                x = O.scheduler_step(O.true_cfg_combine(c, u, 4.0), x, sigmas[i].to(dev), sigmas[i + 1].to(dev))
            err = O.rel_l2(out.float(), x)
        res = {"rel_l2": err, "tol": 1e-2, "ok": bool(err < 1e-2), "steps": steps,
               "model": "tiny MMDiT" if tiny else "2 of 60 blocks at full width (d=3072, 24x128 heads), random-init with biases",
               "shapes": f"batch {B}, {s_img} image + {s_txt} text tokens (the timed workload's)",
               "path": "B200QwenImagePipeline.__call__" + (" -> ParallelDenoiser (same groups layout, exchange, graph capture)" if world > 1 else ""),
               "against": "oracle loop in fp32 on rank 0's GPU, same weights"}
    den.close()
    den.pipe.clear_sessions()
    if world > 1 and rank == 0:
        # the same call on rank 0 alone (no sharding, no exchange): what the multi-GPU layout itself changes
        from vllm_omni_draft_b200.pipeline import B200QwenImagePipeline
        den.tr.set_sequence_parallel(None)
        solo = B200QwenImagePipeline(den.tr)
        solo.use_cuda_graph = use_graph
        one = solo(prompt_embeds=cond, prompt_embeds_mask=mask, negative_prompt_embeds=uncond,
                   negative_prompt_embeds_mask=mask, true_cfg_scale=4.0, num_inference_steps=steps, latents=lat,
                   height=wl["H"], width=wl["W"], output_type="latent").images
        torch.cuda.synchronize()
        # two bf16-operand paths with different tile shapes / summation orders: their roundings are independent, so
        # they sit about as far from each other as sqrt(2) x either one's distance to the fp32 oracle
        res["rel_l2_vs_single_gpu_cuda_path"] = O.rel_l2(out.float(), one.float())
        res["single_gpu_cuda_path_rel_l2"] = O.rel_l2(one.float(), x)
        solo.clear_sessions()
    del den
    torch.cuda.empty_cache()
    return res


SECONDARY_N1 = {   # single-GPU figures the secondary lines are compared with (same bench.py, --workload, 1 GPU)
    "cfg4": ("profiles/r02_bench_cfg4_n1.json", None),
    "cfg5": ("profiles/r02_bench_cfg5_n1.json", None),
}


def _n1_ms(name):
    try:
        with open(os.path.join(ROOT, SECONDARY_N1[name][0])) as fh:
            return float(json.load(fh)["ms_per_step"])
    except (OSError, ValueError, KeyError):
        return None


def time_secondary(name, den, world, rank, dev, barrier, max_over_ranks, steps=6, timed=5):
    """One short run of another BASELINE.json config on the SAME process group and weights (configs[3] = cfg4,
    configs[4] = cfg5: the shapes Ulysses exists for), appended to the cfg-2 line as `secondary`.  Same method as
    the main line: product call from pinned host buffers (setup inside the first call), then `timed` scheduler
    steps between CUDA events, max over ranks, then the product call again for the end-to-end figure."""
    wl = make_workload(name, world)
    H, W, s_txt, B, s_img = wl["H"], wl["W"], wl["s_txt"], wl["batch"], wl["s_img"]
    g1, g2, g3 = (torch.Generator().manual_seed(s) for s in (1, 2, 42))
    cond_h = torch.randn(B, s_txt, D_TXT, generator=g1).bfloat16().pin_memory()
    uncond_h = torch.randn(B, s_txt, D_TXT, generator=g2).bfloat16().pin_memory()
    lat_h = torch.randn(B, s_img, 64, generator=g3).bfloat16().pin_memory()
    mask_h = torch.ones(B, s_txt, dtype=torch.bool)
    out_h = torch.empty_like(lat_h).pin_memory()

    def call():
        out = den.pipe(prompt_embeds=cond_h, prompt_embeds_mask=mask_h, negative_prompt_embeds=uncond_h,
                       negative_prompt_embeds_mask=mask_h, true_cfg_scale=4.0, num_inference_steps=steps,
                       latents=lat_h, height=H, width=W, output_type="latent").images
        out_h.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    barrier()
    t0 = time.perf_counter()
    call()
    t_first = time.perf_counter() - t0
    for i in range(2):
        den.step(i)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(timed):
        den.step(i % steps)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    t0 = time.perf_counter()
    call()
    t_second = time.perf_counter() - t0
    ms, t_first, t_second, bad = max_over_ranks(ms, t_first, t_second, 1.0 if den.exchange_timed_out() else 0.0)
    if bad:
        raise RuntimeError("a peer flag barrier timed out")
    per = ms / timed
    flops = 2 * algorithmic_flops_per_forward(s_img, s_txt, B)
    res = {"workload": wl["config"]["workload"], "name": name, "parallelism": None, "steps": timed, "warmup": 2,
           "ms_per_step": per, "value": 1e3 / per, "unit": "steps/s", "images_per_s": B * 1e3 / per / 50,
           "step_tflops": flops / (per * 1e-3) / 1e12,
           "e2e": {"value": steps / t_second, "unit": "steps/s", "num_inference_steps": steps, "call_s": t_second,
                   "first_call_s": t_first, "h2d_bytes_per_step": (cond_h.numel() + uncond_h.numel() + lat_h.numel()) * 2 / steps,
                   "d2h_bytes_per_step": out_h.numel() * 2 / steps},
           "result_checksum": float(out_h.float().abs().mean())}
    n1 = _n1_ms(name)
    if n1:
        res["n1_ms_per_step"] = n1
        res["n1_source"] = SECONDARY_N1[name][0] + " (round 2, same command at --gpus 1)"
        res["speedup_vs_n1"] = n1 / per
        res["fraction_of_linear"] = n1 / per / world
    return wl, res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=list(WORKLOADS))
    ap.add_argument("--e2e-steps", type=int, default=50, help="num_inference_steps of the timed pipeline call")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernels", action="store_true", help="skip the per-kernel table (multi-config sweeps)")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the short cfg-4 (N >= 4) / cfg-5 (N = 8) runs appended to the cfg-2 line as `secondary`")
    ap.add_argument("--full-step", action="store_true",
                    help="--impl reference only: also time ONE whole cfg-2 step on the host cores (~1-2 min), "
                         "reported as cpu_baseline.full_step beside the extrapolated figure")
    ap.add_argument("--layers", type=int, default=None, help="debug only: fewer blocks (marks the line invalid)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    wl = make_workload(args.workload, args.gpus)
    if args.impl == "reference":
        run_reference_arm(args, wl)
        return

    from vllm_omni_draft_b200 import ops
    from vllm_omni_draft_b200.config import MMDiTConfig
    from vllm_omni_draft_b200.launcher import ParallelDenoiser, init_distributed
    import torch.distributed as dist

    world, rank, local_rank = init_distributed(args.gpus)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    tiny = args.workload == "tiny"
    cfg = MMDiTConfig.tiny() if tiny else MMDiTConfig.qwen_image_20b()
    if args.layers:
        cfg = MMDiTConfig(num_layers=args.layers)
    H, W, s_img, s_txt, B = wl["H"], wl["W"], wl["s_img"], wl["s_txt"], wl["batch"]
    n_sched, K, Wm, E = 50, args.steps, args.warmup, args.e2e_steps
    use_graph = not args.no_graph

    den = ParallelDenoiser(cfg, dev, world, rank, seed=0, use_cuda_graph=use_graph)
    pipe = den.pipe
    pipe.use_cuda_graph = use_graph
    lay = den.layout
    wl["config"]["parallelism"] = (
        "single" if world == 1 else
        (f"cfg{lay.cfg_degree}" if lay.cfg_degree > 1 else "cfg-sequential")
        + (f" x ulysses{lay.sp_degree} ({'fused peer-memory scatter kernels' if den.exchange == 'peer' else 'NCCL all_to_all_single'})"
           if lay.sp_degree > 1 else ""))
    wl["config"]["residual_streams"] = str(den.tr.residual_dtype).replace("torch.", "")
    # synthetic inputs: cond N(0,1) seed 1, uncond seed 2, latents seed 42 — in pinned host memory
    g1, g2, g3 = (torch.Generator().manual_seed(s) for s in (1, 2, 42))
    cond_h = torch.randn(B, s_txt, cfg.joint_attention_dim, generator=g1).bfloat16().pin_memory()
    uncond_h = torch.randn(B, s_txt, cfg.joint_attention_dim, generator=g2).bfloat16().pin_memory()
    lat_h = torch.randn(B, s_img, 64, generator=g3).bfloat16().pin_memory()
    mask_h = torch.ones(B, s_txt, dtype=torch.bool)
    out_h = torch.empty_like(lat_h).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(*vals):
        if world == 1:
            return list(vals)
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t]

    # ---------------- 1. end to end through the product call: host buffers in, host buffer out
    def product_call(steps):
        out = pipe(prompt_embeds=cond_h, prompt_embeds_mask=mask_h, negative_prompt_embeds=uncond_h,
                   negative_prompt_embeds_mask=mask_h, true_cfg_scale=4.0, num_inference_steps=steps, latents=lat_h,
                   height=H, width=W, output_type="latent").images
        out_h.copy_(out, non_blocking=True)               # device -> pinned host
        torch.cuda.current_stream().synchronize()         # the caller holds the result

    barrier()
    t0 = time.perf_counter()
    product_call(E)                                        # first call: session setup inside
    t_first = time.perf_counter() - t0
    first_checksum = float(out_h.float().abs().mean())

    # ---------------- 2. device-resident timing: inputs already in HBM, K scheduler steps
    if world == 1:
        sess = next(iter(pipe._sessions.values()))
        step = sess.step
        n0 = ops.launch_count
        sess._run_step(sess.dts[0:1], sess.mod_table[0:1])                # one eager step, to count its launches
        launches_per_step = ops.launch_count - n0
    else:
        step = den.step
        launches_per_step = den.launches_per_step
    for i in range(Wm):
        step(i % E)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()     # no-op unless run under `ncu --profile-from-start off`
    ev0.record()
    for i in range(K):
        step((Wm + i) % E)
    ev1.record()
    barrier()
    torch.cuda.profiler.stop()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    timed_out = 1.0 if den.exchange_timed_out() else 0.0

    # ---------------- 1b. the product call again: setup is cached now, this is the e2e number
    barrier()
    t0 = time.perf_counter()
    product_call(E)
    t_second = time.perf_counter() - t0
    ms, t_first, t_second, timed_out = max_over_ranks(ms, t_first, t_second, timed_out)
    if timed_out:
        raise RuntimeError("a peer flag barrier timed out during the timed region: the numbers are invalid")
    h2d = (cond_h.numel() + uncond_h.numel() + lat_h.numel()) * 2
    d2h = out_h.numel() * 2
    second_checksum = float(out_h.float().abs().mean())

    # ---------------- 2b. the configs Ulysses exists for, on the same groups and weights (N >= 4, default workload)
    # A failure or a stall in here must not cost the cfg-2 line: errors are recorded per entry, and a watchdog
    # prints the line without `secondary` and ends the process if this part overruns its budget.
    secondary, sec_wl = [], []
    sec_names = [] if (args.no_secondary or args.workload != "cfg2" or tiny or args.layers) else \
        (["cfg4"] if world == 4 else ["cfg4", "cfg5"] if world == 8 else [])
    watchdog = None
    if sec_names:
        import threading
        partial = {"metric": "denoise steps/s", "value": 1e3 / (ms / K), "unit": "steps/s", "n_gpus": world, "steps": K,
                   "warmup": Wm, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong",
                   "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": wl["config"],
                   "e2e": {"value": E / t_second, "unit": "steps/s", "h2d_bytes_per_step": h2d / E,
                           "d2h_bytes_per_step": d2h / E},
                   "gpu_launches": launches_per_step * K, "clocks": clocks,
                   "secondary_error": "the secondary runs overran their time budget; cfg-2 numbers above are complete"}

        def give_up():
            if rank == 0:
                print(json.dumps(partial), flush=True)
            os._exit(0)

        watchdog = threading.Timer(float(os.environ.get("QI_SECONDARY_BUDGET_S", "300")), give_up)
        watchdog.daemon = True
        watchdog.start()
        for name in sec_names:
            try:
                w2, r2 = time_secondary(name, den, world, rank, dev, barrier, max_over_ranks)
                r2["parallelism"] = wl["config"]["parallelism"]
                secondary.append(r2)
                sec_wl.append(w2)
            except Exception as exc:      # noqa: BLE001 — recorded, never fatal for the cfg-2 line
                secondary.append({"name": name, "error": f"{type(exc).__name__}: {exc}"[:300]})
                break                     # the ranks may no longer be in step: no further collective work here

    # ---------------- 3. parity of this layout (every rank takes part)
    den.close()
    pipe.clear_sessions()
    parity = None
    if not args.no_parity:
        parity = layout_parity(world, rank, dev, wl, use_graph, tiny)
        if all("error" not in r for r in secondary):
            for w2, r2 in zip(sec_wl, secondary):
                try:
                    r2["parity"] = layout_parity(world, rank, dev, w2, use_graph, tiny)
                except Exception as exc:  # noqa: BLE001
                    r2["parity"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
                    break
    if watchdog is not None:
        watchdog.cancel()
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):      # every GEMM schedule measured or loaded by this process (feeds tuned_schedules.json)
        with open(os.path.join(out_dir, f"tuned_schedules_n{world}.json"), "w") as f:
            json.dump(ops.tuned_schedules(), f, indent=0, sort_keys=True)
    burst, sustained, hbm, src = load_peaks()
    ms_step = ms / K
    flops_step = 2 * algorithmic_flops_per_forward(s_img, s_txt, batch=B, layers=cfg.num_layers, d=cfg.inner_dim,
                                                   d_txt=cfg.joint_attention_dim)
    tf_step = flops_step / (ms_step * 1e-3) / 1e12
    line = {"metric": "denoise steps/s", "value": 1e3 / ms_step, "unit": "steps/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": wl["config"],
            "s_per_image": t_second * (50.0 / E),
            "s_per_image_note": f"wall clock of one {E}-step pipeline call (host in, host out), second call"
                                + ("" if E == 50 else f", scaled x{50.0 / E:.2f} to 50 steps"),
            "images_per_s": B / (t_second * (50.0 / E)),
            "step_tflops": tf_step,
            "step_frac_of_measured_burst_peak": tf_step / (burst * world),
            "step_frac_of_measured_sustained_peak": tf_step / (sustained * world),
            "step_frac_of_nominal_peak": tf_step / (2250.0 * world),
            "e2e": {"value": E / t_second, "unit": "steps/s", "h2d_bytes_per_step": h2d / E, "d2h_bytes_per_step": d2h / E,
                    "call": "B200QwenImagePipeline.__call__(prompt_embeds=, negative_prompt_embeds=, latents=<pinned host>, "
                            f"num_inference_steps={E}, output_type='latent') -> pinned host",
                    "num_inference_steps": E, "call_s": t_second, "first_call_s": t_first,
                    "setup_ms": (t_first - t_second) * 1e3, "h2d_bytes_per_call": h2d, "d2h_bytes_per_call": d2h,
                    "result_checksums": [first_checksum, second_checksum]},
            "gpu_launches": launches_per_step * K, "gpu_launches_per_step": launches_per_step,
            "clocks": clocks, "cuda_graph": use_graph, "parity": parity}
    if secondary:
        line["secondary"] = secondary
    if args.layers:
        line["invalid"] = "debug run with --layers"
    if parity is not None and not parity["ok"]:
        line["invalid"] = f"layout parity failed: rel-L2 {parity['rel_l2']:.3e} > {parity['tol']}"
    # ---------------- 4. roofline of the dominant kernel + per-kernel table (rank 0, single GPU work)
    del den
    torch.cuda.empty_cache()
    if not args.no_kernels and not tiny:
        from vllm_omni_draft_b200.roofline import kernel_rooflines
        line.update(kernel_rooflines(burst, sustained, hbm, src))
        traffic, kname = committed_traffic()
        r = line["roofline"]
        r["traffic"] = traffic
        r["traffic_source"] = f"dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full, {os.path.relpath(TRAFFIC_CSV, ROOT)} ({kname[:60]})"
        # algorithmic bytes of that launch: A (batch 2) + W (img + txt) + C
        r["algorithmic_bytes"] = 2.0 * 2 * (4096 + 512) * D + 2 * 2.0 * 4 * D * D + 2.0 * 2 * (4096 + 512) * 4 * D
    # ---------------- 5. the reference's CPU path on this box's host cores
    if world == 1 and not args.no_cpu_baseline and not tiny:
        cb, _ = cpu_baseline_block(s_img, s_txt, B, 50, repeats=3, warm=1)
        cb["cfg1_loop"] = cpu_cfg1_loop()
        line["cpu_baseline"] = cb
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
