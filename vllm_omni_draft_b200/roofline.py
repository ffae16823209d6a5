"""Per-kernel timing at the cfg-2 shapes with CUDA events (bench.py's `roofline` / `kernels` keys).

Every kernel is captured into a CUDA graph of `reps` back-to-back launches and the graph is timed,
so the number is the device time of the kernel and not the Python launch rate.  GEMM launches
rotate through `nw` weight sets (> L2 in total) so operands are not L2-resident; the element-wise
kernels work on activations that the previous kernel of the block leaves in L2, which is also how
they run inside a step (noted per entry)."""
from __future__ import annotations

import torch

from . import ops, rope


def _time_graph(fn, reps: int) -> float:
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(reps):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


def kernel_rooflines(burst_tf: float, sustained_tf: float, hbm_gbs: float, src: str, d=3072, H=24, Dh=128,
                     s_img=4096, s_txt=512, nw=8, reps=24) -> dict:
    dev = "cuda"
    S = s_img + s_txt
    g = torch.Generator(device=dev).manual_seed(0)
    r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).bfloat16()
    emp = lambda *s: torch.empty(*s, device=dev, dtype=torch.bfloat16)
    xi, xt = r(1, s_img, d), r(1, s_txt, d)
    w_up, b_up = [r(2, 4 * d, d, sc=0.02) for _ in range(nw)], r(2, 4 * d)
    mi, mt = emp(1, s_img, 4 * d), emp(1, s_txt, 4 * d)
    kern = {}
    fl_mlp = 2.0 * S * 4 * d * d
    t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(xi, w_up[i % nw][0], b_up[0], mi, ops.EPI_BIAS_GELU_TANH),
                                        ops.gemm_problem(xt, w_up[i % nw][1], b_up[1], mt, ops.EPI_BIAS_GELU_TANH)]), reps)
    roof = {"bound": "tensor", "kernel": "gemm_pair_bf16_kernel<256> (MLP-up + bias + GELU-tanh, img+txt streams in one launch, cta_group::2)",
            "achieved": fl_mlp / t / 1e12, "peak": sustained_tf, "unit": "TFLOP/s", "frac": fl_mlp / t / 1e12 / sustained_tf,
            "traffic": None, "peak_source": f"{src}: cuBLAS bf16 sustained (kernel timed in a back-to-back loop)",
            "launch_us": t * 1e6, "flops_per_launch": fl_mlp, "frac_of_burst": fl_mlp / t / 1e12 / burst_tf,
            "frac_of_nominal_2250": fl_mlp / t / 1e12 / 2250.0}
    w_dn = [w.view(2, d, 4 * d) for w in w_up]
    gate, ho, eo = r(1, d), emp(1, s_img, d), emp(1, s_txt, d)
    t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(mi, w_dn[i % nw][0], b_up[0, :d], ho, 2, xi, gate),
                                        ops.gemm_problem(mt, w_dn[i % nw][1], b_up[1, :d], eo, 2, xt, gate)]), reps)
    kern["gemm_mlp_down_gated_residual"] = {"us": t * 1e6, "tflops": fl_mlp / t / 1e12}
    w_qkv = [w.view(2, 4 * d, d)[:, :3 * d].contiguous() for w in w_up[:4]]
    qkv = emp(1, S, 3 * d)
    fr = rope.joint_cos_sin((16, 56, 56), 10000.0, [(1, int(s_img ** 0.5), int(s_img ** 0.5))], s_txt).to(dev)
    nwt = r(4, Dh) + 1
    qk = dict(epilogue=ops.EPI_QKV_NORM_ROPE, freqs=fr, heads=H, head_dim=Dh)
    t = _time_graph(lambda i: ops.gemm([
        ops.gemm_problem(xi, w_qkv[i % 4][0], b_up[0, :3 * d], qkv[:, s_txt:], qk_norm=(nwt[0], nwt[1]), rope_pos0=s_txt, **qk),
        ops.gemm_problem(xt, w_qkv[i % 4][1], b_up[1, :3 * d], qkv[:, :s_txt], qk_norm=(nwt[2], nwt[3]), rope_pos0=0, **qk)]), reps)
    kern["gemm_qkv_qknorm_rope"] = {"us": t * 1e6, "tflops": 2.0 * S * 3 * d * d / t / 1e12,
                                    "note": "QK-RMSNorm + MSRoPE fused into the QKV GEMM epilogue"}
    w_o = [w.view(2, 4 * d, d)[:, :d].contiguous() for w in w_up[:4]]
    att = r(1, S, d)
    t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(att[:, s_txt:], w_o[i % 4][0], b_up[0, :d], ho, 2, xi, gate),
                                        ops.gemm_problem(att[:, :s_txt], w_o[i % 4][1], b_up[1, :d], eo, 2, xt, gate)]), reps)
    kern["gemm_out_proj_gated_residual"] = {"us": t * 1e6, "tflops": 2.0 * S * d * d / t / 1e12}
    qkv.copy_(r(1, S, 3 * d))
    q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
    o = emp(1, S, H, Dh)
    t = _time_graph(lambda i: ops.attention(q, k, v, o), reps)
    fa = 4.0 * S * S * d
    kern["attention_fwd"] = {"us": t * 1e6, "tflops": fa / t / 1e12, "frac_of_sustained": fa / t / 1e12 / sustained_tf,
                             "frac_of_nominal_2250": fa / t / 1e12 / 2250.0}
    mod, out = r(1, 6 * d, sc=0.1), emp(s_img, d)
    t = _time_graph(lambda i: ops.adaln_modulate(xi.view(s_img, d), mod[:, :d], mod[:, d:2 * d], out, s_img), reps)
    by = 4.0 * s_img * d
    kern["adaln_modulate_img"] = {"us": t * 1e6, "gbs": by / t / 1e9, "frac_of_hbm": by / t / 1e9 / hbm_gbs,
                                  "bytes": by, "note": "50 MB working set, L2-resident as inside a step"}
    q2 = qkv.view(S, 3 * d)
    t = _time_graph(lambda i: ops.qknorm_rope(q2, H, Dh, S, s_txt, nwt[0], nwt[1], nwt[2], nwt[3], fr), reps)
    by = 8.0 * S * d
    kern["qknorm_rope_standalone"] = {"us": t * 1e6, "gbs": by / t / 1e9, "frac_of_hbm": by / t / 1e9 / hbm_gbs, "bytes": by,
                                      "note": "C-ABI entry point kept; inside the step this work runs in the QKV GEMM epilogue"}
    lat = [r(1, s_img, 64) for _ in range(4)]
    t = _time_graph(lambda i: ops.cfg_euler_step(lat[0], lat[1], lat[2], lat[3], 4.0, dt=-0.02), reps)
    by = 4.0 * s_img * 64 * 2
    kern["cfg_euler_step"] = {"us": t * 1e6, "gbs": by / t / 1e9, "bytes": by, "note": "2.1 MB: launch-latency bound"}
    blk = (kern["gemm_qkv_qknorm_rope"]["us"] + kern["gemm_out_proj_gated_residual"]["us"] + roof["launch_us"]
           + kern["gemm_mlp_down_gated_residual"]["us"] + kern["attention_fwd"]["us"]
           + 2 * kern["adaln_modulate_img"]["us"] * (S / s_img))
    kern["block_sum_us"] = blk
    kern["step_from_kernel_sum_ms"] = blk * 120 / 1e3
    return {"roofline": roof, "kernels": kern}
