"""Per-kernel timing at the cfg-2 shapes with CUDA events (bench.py's `roofline` / `kernels` keys).

Every kernel is captured into a CUDA graph of `reps` back-to-back launches and the graph is timed, so the
number is the device time of the kernel and not the Python launch rate.  The launches are the ones a step
issues: batch 2 (cond and uncond stacked), img + txt stream in one launch.  GEMM launches rotate through `nw`
weight sets (> L2 in total) so operands are not L2-resident.  The memory-bound kernels (K1, K2, K5) are timed
twice: over a rotating set of buffers larger than L2 (`gbs` = true HBM rate, the roofline number) and on one
L2-resident buffer (`gbs_l2_resident`, how they meet their input inside a step, right after its producer).

Library bars measured beside our kernels under the identical graph: cuBLASLt via `F.linear` at
the same shapes (`cublas_same_shape_tflops`) and `F.scaled_dot_product_attention` (cuDNN / flash backends) at
the same (B, H, S, Dh) (`sdpa_same_shape_tflops`).  Library calls appear here only as the comparison."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import ops, rope

L2_BYTES = 126e6


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
    best = None
    for _ in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / reps * 1e-3
        best = t if best is None else min(best, t)
    return best


def _sdpa_bars(q, k, v, reps: int, flops: float) -> dict:
    """F.scaled_dot_product_attention at the same shape, per backend that runs here."""
    out = {}
    try:
        from torch.nn.attention import SDPBackend, sdpa_kernel
    except Exception as e:  # pragma: no cover
        return {"error": repr(e)}
    qh, kh, vh = (t.permute(0, 2, 1, 3).contiguous() for t in (q, k, v))        # [B, H, S, Dh]
    for name in ("CUDNN_ATTENTION", "FLASH_ATTENTION"):
        be = getattr(SDPBackend, name, None)
        if be is None:
            continue
        try:
            with sdpa_kernel([be]):
                t = _time_graph(lambda i: F.scaled_dot_product_attention(qh, kh, vh), reps)
            out[name.lower()] = {"us": t * 1e6, "tflops": flops / t / 1e12}
        except Exception as e:
            out[name.lower()] = {"error": str(e).splitlines()[0][:160]}
    return out


def kernel_rooflines(burst_tf: float, sustained_tf: float, hbm_gbs: float, src: str, d=3072, H=24, Dh=128,
                     s_img=4096, s_txt=512, nw=8, reps=24, batch=2, library_bars: bool = True) -> dict:
    dev = "cuda"
    S, B = s_img + s_txt, batch
    g = torch.Generator(device=dev).manual_seed(0)
    r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).bfloat16()
    emp = lambda *s: torch.empty(*s, device=dev, dtype=torch.bfloat16)
    xi, xt = r(B, s_img, d), r(B, s_txt, d)
    w_up, b_up = [r(2, 4 * d, d, sc=0.02) for _ in range(nw)], r(2, 4 * d)
    mi, mt = emp(B, s_img, 4 * d), emp(B, s_txt, 4 * d)
    kern = {}

    def lib_pair(a_i, a_t, ws, cols):
        """cuBLASLt through F.linear on the same two problems (img, txt) inside the same kind of graph."""
        if not library_bars:
            return None
        a_i, a_t = a_i.reshape(-1, a_i.shape[-1]), a_t.reshape(-1, a_t.shape[-1])
        return _time_graph(lambda i: (F.linear(a_i, ws[i % len(ws)][0][:cols]), F.linear(a_t, ws[i % len(ws)][1][:cols])), reps)

    # ---- MLP-up (dominant launch): bias + GELU-tanh
    fl_mlp = 2.0 * B * S * 4 * d * d
    t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(xi, w_up[i % nw][0], b_up[0], mi, ops.EPI_BIAS_GELU_TANH),
                                        ops.gemm_problem(xt, w_up[i % nw][1], b_up[1], mt, ops.EPI_BIAS_GELU_TANH)]), reps)
    ach = fl_mlp / t / 1e12
    roof = {"bound": "tensor",
            "kernel": f"gemm_pair_bf16_kernel (MLP-up + bias + GELU-tanh, img+txt streams in one launch, batch {B} as in the step, cta_group::2)",
            "achieved": ach, "peak": burst_tf, "unit": "TFLOP/s", "frac": ach / burst_tf, "traffic": None,
            "peak_source": f"{src}: cuBLAS bf16 burst (the kernel is timed alone, {reps} back-to-back launches)",
            "launch_us": t * 1e6, "flops_per_launch": fl_mlp, "frac_of_sustained": ach / sustained_tf,
            "frac_of_nominal_2250": ach / 2250.0}
    tl = lib_pair(xi, xt, w_up, 4 * d)
    if tl:
        roof["cublas_same_shape_tflops"] = fl_mlp / tl / 1e12
        roof["cublas_note"] = "F.linear (cuBLASLt), no bias/GELU epilogue, two calls (img, txt)"
    # ---- MLP-down + gated residual
    w_dn = [w.view(2, d, 4 * d) for w in w_up]
    # gated-residual launches as the step issues them: fp32 residual streams updated in place, fp32 gate row
    r32 = lambda *s, sc=1.0: torch.randn(*s, device=dev, generator=g) * sc
    gate, ho, eo = r32(1, d, sc=0.1).expand(B, d), r32(B, s_img, d), r32(B, s_txt, d)
    t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(mi, w_dn[i % nw][0], b_up[0, :d], ho, 2, ho, gate),
                                        ops.gemm_problem(mt, w_dn[i % nw][1], b_up[1, :d], eo, 2, eo, gate)]), reps)
    kern["gemm_mlp_down_gated_residual"] = {"us": t * 1e6, "tflops": fl_mlp / t / 1e12}
    tl = lib_pair(mi, mt, w_dn, d)
    if tl:
        kern["gemm_mlp_down_gated_residual"]["cublas_same_shape_tflops"] = fl_mlp / tl / 1e12
    # ---- QKV + QK-RMSNorm + MSRoPE
    w_qkv = [w.view(2, 4 * d, d)[:, :3 * d].contiguous() for w in w_up[:4]]
    qkv = emp(B, S, 3 * d)
    fr = rope.joint_cos_sin((16, 56, 56), 10000.0, [(1, int(s_img ** 0.5), int(s_img ** 0.5))], s_txt).to(dev)
    nwt = r(4, Dh) + 1
    qk = dict(epilogue=ops.EPI_QKV_NORM_ROPE, freqs=fr, heads=H, head_dim=Dh)
    t = _time_graph(lambda i: ops.gemm([
        ops.gemm_problem(xi, w_qkv[i % 4][0], b_up[0, :3 * d], qkv[:, s_txt:], qk_norm=(nwt[0], nwt[1]), rope_pos0=s_txt, **qk),
        ops.gemm_problem(xt, w_qkv[i % 4][1], b_up[1, :3 * d], qkv[:, :s_txt], qk_norm=(nwt[2], nwt[3]), rope_pos0=0, **qk)]), reps)
    fl = 2.0 * B * S * 3 * d * d
    kern["gemm_qkv_qknorm_rope"] = {"us": t * 1e6, "tflops": fl / t / 1e12,
                                    "note": "QK-RMSNorm + MSRoPE fused into the QKV GEMM epilogue"}
    tl = lib_pair(xi, xt, w_qkv, 3 * d)
    if tl:
        kern["gemm_qkv_qknorm_rope"]["cublas_same_shape_tflops"] = fl / tl / 1e12
    # ---- out-projection + gated residual
    w_o = [w.view(2, 4 * d, d)[:, :d].contiguous() for w in w_up[:4]]
    att = r(B, S, d)
    t = _time_graph(lambda i: ops.gemm([ops.gemm_problem(att[:, s_txt:], w_o[i % 4][0], b_up[0, :d], ho, 2, ho, gate),
                                        ops.gemm_problem(att[:, :s_txt], w_o[i % 4][1], b_up[1, :d], eo, 2, eo, gate)]), reps)
    fl = 2.0 * B * S * d * d
    kern["gemm_out_proj_gated_residual"] = {"us": t * 1e6, "tflops": fl / t / 1e12}
    tl = lib_pair(att[:, s_txt:].contiguous(), att[:, :s_txt].contiguous(), w_o, d)
    if tl:
        kern["gemm_out_proj_gated_residual"]["cublas_same_shape_tflops"] = fl / tl / 1e12
    # ---- attention
    qkv.copy_(r(B, S, 3 * d))
    q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
    o = emp(B, S, H, Dh)
    t = _time_graph(lambda i: ops.attention(q, k, v, o), reps)
    fa = 4.0 * B * S * S * d
    kern["attention_fwd"] = {"us": t * 1e6, "tflops": fa / t / 1e12, "frac_of_burst": fa / t / 1e12 / burst_tf,
                             "frac_of_sustained": fa / t / 1e12 / sustained_tf,
                             "frac_of_nominal_2250": fa / t / 1e12 / 2250.0}
    if library_bars:
        kern["attention_fwd"]["sdpa_same_shape_tflops"] = _sdpa_bars(q, k, v, reps, fa)
    del w_up, w_dn, w_qkv, w_o, mi, mt
    torch.cuda.empty_cache()

    # ---- memory-bound kernels: rotating buffers larger than L2 (true HBM rate) and one resident buffer
    def hbm_entry(name, fn, by, nbuf, note):
        t_hbm = _time_graph(lambda i: fn(i % nbuf), reps)
        t_l2 = _time_graph(lambda i: fn(0), reps)
        kern[name] = {"us": t_hbm * 1e6, "gbs": by / t_hbm / 1e9, "frac_of_hbm": by / t_hbm / 1e9 / hbm_gbs,
                      "bytes": by, "working_set_bytes": by * nbuf, "us_l2_resident": t_l2 * 1e6,
                      "gbs_l2_resident": by / t_l2 / 1e9, "note": note}

    rows = B * s_img
    nb = max(3, int(2.5 * L2_BYTES // (6.0 * rows * d)) + 1)
    xs, ys = [r32(rows, d) for _ in range(nb)], [emp(rows, d) for _ in range(nb)]
    mod = r32(1, 6 * d, sc=0.1)
    sh, sc_ = mod[:, :d].expand(B, d), mod[:, d:2 * d].expand(B, d)
    hbm_entry("adaln_modulate_img", lambda j: ops.adaln_modulate_pair([(xs[j], sh, sc_, ys[j], s_img)]),
              6.0 * rows * d, nb, f"K1 as in the step: fp32 residual rows in, bf16 out, fp32 modulation; batch {B}; {nb} rotating in/out pairs")
    del xs, ys
    nb = max(3, int(2.5 * L2_BYTES // (8.0 * S * d)) + 1)
    qs = [r(S, 3 * d) for _ in range(nb)]
    hbm_entry("qknorm_rope_standalone",
              lambda j: ops.qknorm_rope(qs[j], H, Dh, S, s_txt, nwt[0], nwt[1], nwt[2], nwt[3], fr),
              8.0 * S * d, nb, "K2 C-ABI entry point (batch 1); inside the step this work runs in the QKV GEMM epilogue")
    del qs
    nb = 48
    lat = [[r32(1, s_img, 64) for _ in range(3)] + [emp(1, s_img, 64)] for _ in range(nb)]
    hbm_entry("cfg_euler_step", lambda j: ops.cfg_euler_step_f32(lat[j][0], lat[j][1], lat[j][2], lat[j][2], lat[j][3], 4.0, dt=-0.02),
              (4 * 4.0 + 2.0) * s_img * 64, nb, "K5 in fp32 (master latents in place + bf16 copy), 4.7 MB per launch: launch-latency bound")
    blk = (kern["gemm_qkv_qknorm_rope"]["us"] + kern["gemm_out_proj_gated_residual"]["us"] + roof["launch_us"]
           + kern["gemm_mlp_down_gated_residual"]["us"] + kern["attention_fwd"]["us"]
           + 2 * kern["adaln_modulate_img"]["us_l2_resident"] * (S / s_img))
    kern["block_sum_us"] = blk
    kern["step_from_kernel_sum_ms"] = blk * 60 / 1e3 * (2 / B)
    # `roofline` describes the dominant kernel (the largest share of the step); `roofline_worst` is the tensor-bound
    # kernel furthest below the same peak, i.e. the one that decides whether the step meets its target
    # (the headline object must not flatter).  Same peak, same timing method as `roofline`.
    tensor = {n: kv for n, kv in kern.items() if isinstance(kv, dict) and "tflops" in kv}
    tensor["gemm_mlp_up_gelu"] = {"us": roof["launch_us"], "tflops": roof["achieved"]}
    wname = min(tensor, key=lambda n: tensor[n]["tflops"])
    w = tensor[wname]
    worst = {"bound": "tensor", "kernel": wname, "achieved": w["tflops"], "peak": burst_tf, "unit": "TFLOP/s",
             "frac": w["tflops"] / burst_tf, "frac_of_sustained": w["tflops"] / sustained_tf,
             "frac_of_nominal_2250": w["tflops"] / 2250.0, "launch_us": w["us"],
             "share_of_block": w["us"] / blk, "peak_source": roof.get("peak_source"), "traffic": None}
    if "sdpa_same_shape_tflops" in w:
        worst["library_same_shape_tflops"] = w["sdpa_same_shape_tflops"]
    elif "cublas_same_shape_tflops" in w:
        worst["library_same_shape_tflops"] = w["cublas_same_shape_tflops"]
    return {"roofline": roof, "roofline_worst": worst, "kernels": kern}
