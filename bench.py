#!/usr/bin/env python
"""bench.py — denoise steps/s of the Qwen-Image hot path (BASELINE.json metric) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path (oracle)

One "step" = everything between two successive latents with true CFG: cond forward + uncond forward
of the 60-block MMDiT + the fused CFG/Euler kernel.  Workload at N=1: BASELINE.json configs[1]
(Qwen-Image 20B, 1024x1024 -> 4096 image tokens, 512 text tokens, bf16, random-init, synthetic
prompt embeddings and latents).  N=2: CFG-parallel (cond / uncond ranks); N=4/8: CFG-parallel x
Ulysses sequence parallel (strong scaling of the same job).  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

D, L, HEADS, DH, D_TXT = 3072, 60, 24, 128, 3584


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
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


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
    img_f, txt_f = O.rope_freqs(cfg, [(1, int(s_img ** 0.5), int(s_img ** 0.5))], s_txt)
    times = []
    with torch.no_grad():
        for _ in range(repeats):
            t0 = time.perf_counter()
            hh, ee = h, e
            for i in range(n_blocks):
                ee, hh = O.block_forward(sd, i, cfg, hh, ee, temb, img_f, txt_f)
            times.append(time.perf_counter() - t0)
    return times, torch.get_num_threads()


def run_reference_arm(args, workload):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    s_img, s_txt = workload["s_img"], workload["s_txt"]
    n_blocks = 2
    times, threads = cpu_reference_sample(s_img, s_txt, n_blocks, repeats=args.warmup + args.steps)
    timed = times[args.warmup:]
    t_sample = sum(timed) / len(timed)
    step_s = t_sample * (L / n_blocks) * 2
    value = 1.0 / step_s
    sample = f"{n_blocks} of {L} MMDiT blocks of one cond forward at full cfg-2 shapes, fp32, extrapolated x{L // n_blocks} x2 forwards"
    line = {"impl": "reference", "metric": "denoise steps/s", "value": value, "unit": "steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload["config"],
            "cpu_baseline": {"value": value, "unit": "steps/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "s_per_image": step_s * workload["num_inference_steps"], "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "tiny", "cfg4"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--layers", type=int, default=None, help="debug only: fewer blocks (marks the line invalid)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3

    if args.workload == "cfg2":
        H = W = 1024
    elif args.workload == "cfg4":
        H = W = 2048
    else:
        H = W = 256
    s_img = (H // 16) * (W // 16)
    s_txt = 512 if args.workload != "tiny" else 32
    workload = {"s_img": s_img, "s_txt": s_txt, "num_inference_steps": 50,
                "config": {"workload": f"Qwen-Image 20B MMDiT (60 blocks, d=3072, 24x128 heads) random-init, "
                                       f"{H}x{W} ({s_img} image tokens), {s_txt} text tokens, batch 1, true CFG "
                                       f"(2 forwards/step), 50-step FlowMatch schedule"
                           if args.workload != "tiny" else "tiny MMDiT (2 blocks, d=256) 256x256",
                           "parallelism": {1: "single", 2: "cfg2", 4: "cfg2 x ulysses2", 8: "cfg2 x ulysses4"}.get(args.gpus, "?"),
                           "l2": "weights of 60 distinct blocks (40.9 GB) stream through every step: inputs larger than L2, no flush needed"}}
    if args.impl == "reference":
        run_reference_arm(args, workload)
        return

    from vllm_omni_draft_b200 import ops
    from vllm_omni_draft_b200.config import MMDiTConfig
    from vllm_omni_draft_b200.launcher import ParallelDenoiser, init_distributed

    world, rank, local_rank = init_distributed(args.gpus)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    cfg = MMDiTConfig.qwen_image_20b() if args.workload != "tiny" else MMDiTConfig.tiny()
    if args.layers:
        cfg = MMDiTConfig(num_layers=args.layers)
    n_steps_sched = 50
    K, Wm = args.steps, args.warmup

    den = ParallelDenoiser(cfg, dev, world, rank, seed=0, use_cuda_graph=not args.no_graph)
    lay = den.layout
    workload["config"]["parallelism"] = (
        "single" if world == 1 else
        (f"cfg{lay.cfg_degree}" if lay.cfg_degree > 1 else "cfg-sequential")
        + (f" x ulysses{lay.sp_degree} ({'fused peer-memory scatter kernels' if den.exchange == 'peer' else 'NCCL all_to_all_single'})"
           if lay.sp_degree > 1 else ""))
    # synthetic inputs: cond N(0,1) seed 1, uncond seed 2, latents seed 42
    g1, g2, g3 = (torch.Generator().manual_seed(s) for s in (1, 2, 42))
    cond_h = torch.randn(1, s_txt, cfg.joint_attention_dim, generator=g1).bfloat16().pin_memory()
    uncond_h = torch.randn(1, s_txt, cfg.joint_attention_dim, generator=g2).bfloat16().pin_memory()
    lat_h = torch.randn(1, s_img, 64, generator=g3).bfloat16().pin_memory()
    out_h = torch.empty_like(lat_h).pin_memory()
    den.setup(cond_h, uncond_h, lat_h, H, W, n_steps_sched, true_cfg_scale=4.0)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident timing: inputs already in HBM
    for i in range(Wm):
        den.step(i % n_steps_sched)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ops.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()     # no-op unless run under `ncu --profile-from-start off`
    ev0.record()
    for i in range(K):
        den.step((Wm + i) % n_steps_sched)
    ev1.record()
    barrier()
    torch.cuda.profiler.stop()
    ms = ev0.elapsed_time(ev1)
    launches = den.launches_per_step * K if den.launches_per_step else ops.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None

    # ---------------- end to end: host buffers in, host buffer out, every step
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        den.load_inputs(cond_h, uncond_h, lat_h)          # pinned host -> device
        den.step((Wm + i) % n_steps_sched)
        den.read_latents(out_h)                            # device -> pinned host
        torch.cuda.current_stream().synchronize()          # the caller sees the result
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])
    h2d = cond_h.numel() * 2 + uncond_h.numel() * 2 + lat_h.numel() * 2
    d2h = out_h.numel() * 2

    if rank != 0:
        den.close()
        if world > 1:
            torch.distributed.destroy_process_group()
        return

    burst, sustained, hbm, src = load_peaks()
    ms_step = ms / K
    flops_step = 2 * algorithmic_flops_per_forward(s_img, s_txt, layers=cfg.num_layers, d=cfg.inner_dim,
                                                   d_txt=cfg.joint_attention_dim)
    line = {"metric": "denoise steps/s", "value": 1e3 / ms_step, "unit": "steps/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload["config"],
            "s_per_image": ms_step * 50 / 1e3,
            "step_tflops": flops_step / (ms_step * 1e-3) / 1e12 ,
            "step_frac_of_measured_sustained_peak": flops_step / (ms_step * 1e-3) / 1e12 / (sustained * world),
            "step_frac_of_nominal_peak": flops_step / (ms_step * 1e-3) / 1e12 / (2250.0 * world),
            "e2e": {"value": 1e3 / (ms_e2e / K), "unit": "steps/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h},
            "gpu_launches": launches, "clocks": clocks, "cuda_graph": not args.no_graph}
    if args.layers:
        line["invalid"] = "debug run with --layers"
    # ---------------- roofline of the dominant kernel + per-kernel table (rank 0, single GPU work)
    den.close()
    del den
    torch.cuda.empty_cache()
    from vllm_omni_draft_b200.roofline import kernel_rooflines
    line.update(kernel_rooflines(burst, sustained, hbm, src))
    # DRAM traffic of the dominant kernel per launch, from the committed `ncu --set full` capture of the same
    # launch (profiles/r01c_ncu_full_summary.csv rows 1-3: dram__bytes_read.sum + dram__bytes_write.sum);
    # algorithmic bytes = A 28.3 MB + W(img+txt) 151.0 MB + C 113.2 MB = 292.5 MB
    line["roofline"]["traffic"] = 287.6e6
    line["roofline"]["traffic_unit"] = "bytes per launch (ncu, profiles/r01c_ncu_full_summary.csv)"
    line["roofline"]["algorithmic_bytes"] = 2.0 * (s_img + s_txt) * D + 2 * 2.0 * 4 * D * D + 2.0 * (s_img + s_txt) * 4 * D
    if world == 1 and not args.no_cpu_baseline and args.workload == "cfg2":
        times, threads = cpu_reference_sample(s_img, s_txt, 2, repeats=2)
        t_sample = min(times)
        step_s = t_sample * (L / 2) * 2
        line["cpu_baseline"] = {"value": 1.0 / step_s, "unit": "steps/s", "cores": threads, "kind": "port",
                                "sample": "2 of 60 MMDiT blocks of one cond forward at full cfg-2 shapes, fp32 oracle, "
                                          "best of 2, extrapolated x30 x2 forwards"}
    print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
