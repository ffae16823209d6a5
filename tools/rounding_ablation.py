#!/usr/bin/env python
"""Which bf16 rounding points of the CUDA path cost how much accuracy against the fp32 oracle?  (CPU, minutes.)

A narrow (d = 512, 4 x 128 heads) but full-DEPTH (60 blocks) random-init MMDiT is run in fp32 with a bf16 rounding
injected at each place the CUDA path stores bf16 — individually switchable — and compared with the plain fp32
oracle forward.  Rounding error per element does not depend on the width, the propagation through depth does, so
depth is kept and width is cut.  Output: relative-L2 of `.sample` with every rounding on, and with each one
removed in turn.  Used to decide where fp32 storage pays."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from oracle import qwen_image_oracle as O  # noqa: E402

POINTS = ["res", "mod", "temb", "xn", "qk_multi", "qkv", "p", "attn", "hid", "out"]


def r(x, on):
    return x.bfloat16().float() if on else x


def forward(sd, cfg, x, enc, t, shapes, s_txt, on):
    H, Dh = cfg.num_attention_heads, cfg.attention_head_dim
    h = F.linear(x, sd["img_in.weight"], sd["img_in.bias"])
    e = F.linear(r(O.rms_norm(enc, sd["txt_norm.weight"]), True), sd["txt_in.weight"], sd["txt_in.bias"])
    h, e = r(h, on["res"]), r(e, on["res"])
    p = "time_text_embed.timestep_embedder."
    sin = r(O.timestep_sinusoid(t), True)
    t1 = r(F.silu(F.linear(sin, sd[p + "linear_1.weight"], sd[p + "linear_1.bias"])), on["temb"])
    temb = r(F.linear(t1, sd[p + "linear_2.weight"], sd[p + "linear_2.bias"]), on["temb"])
    st = r(F.silu(temb), True)           # A operand of the modulation GEMM: bf16 either way
    fi, ft = O.rope_freqs(cfg, shapes, s_txt)
    fi, ft = fi.to(x.device), ft.to(x.device)

    def qk(xx, w, fr):
        if on["qk_multi"]:               # upstream's eager rounding order, as the QKV epilogue reproduces it
            xx = r(xx, True)
            var = xx.pow(2).mean(-1, keepdim=True)
            xx = r(r(xx * torch.rsqrt(var + 1e-6), True) * w, True)
        else:
            var = xx.pow(2).mean(-1, keepdim=True)
            xx = xx * torch.rsqrt(var + 1e-6) * w
        return r(O.apply_rope(xx, fr), on["qkv"])

    for i in range(cfg.num_layers):
        b = f"transformer_blocks.{i}."
        im = r(F.linear(st, sd[b + "img_mod.1.weight"], sd[b + "img_mod.1.bias"]), on["mod"])
        tm = r(F.linear(st, sd[b + "txt_mod.1.weight"], sd[b + "txt_mod.1.bias"]), on["mod"])
        (i1, i2), (t1m, t2m) = im.chunk(2, -1), tm.chunk(2, -1)
        xi, gi1 = O._modulate(O.layer_norm(h), i1)
        xt, gt1 = O._modulate(O.layer_norm(e), t1m)
        xi, xt = r(xi, on["xn"]), r(xt, on["xn"])
        a = b + "attn."
        pr = lambda z, n: F.linear(z, sd[a + n + ".weight"], sd[a + n + ".bias"]).unflatten(-1, (H, Dh))
        iq, ik = qk(pr(xi, "to_q"), sd[a + "norm_q.weight"], fi), qk(pr(xi, "to_k"), sd[a + "norm_k.weight"], fi)
        tq, tk = qk(pr(xt, "add_q_proj"), sd[a + "norm_added_q.weight"], ft), qk(pr(xt, "add_k_proj"), sd[a + "norm_added_k.weight"], ft)
        iv, tv = r(pr(xi, "to_v"), on["qkv"]), r(pr(xt, "add_v_proj"), on["qkv"])
        q = torch.cat([tq, iq], 1).transpose(1, 2)
        k = torch.cat([tk, ik], 1).transpose(1, 2)
        v = torch.cat([tv, iv], 1).transpose(1, 2)
        s = (q @ k.transpose(-1, -2)) * Dh ** -0.5
        pm = torch.exp(s - s.amax(-1, keepdim=True))
        o = (r(pm, on["p"]) @ v) / pm.sum(-1, keepdim=True)
        o = r(o.transpose(1, 2).flatten(2), on["attn"])
        ao_t, ao_i = o[:, :s_txt], o[:, s_txt:]
        h = r(h + gi1 * F.linear(ao_i, sd[a + "to_out.0.weight"], sd[a + "to_out.0.bias"]), on["res"])
        e = r(e + gt1 * F.linear(ao_t, sd[a + "to_add_out.weight"], sd[a + "to_add_out.bias"]), on["res"])
        xi2, gi2 = O._modulate(O.layer_norm(h), i2)
        xt2, gt2 = O._modulate(O.layer_norm(e), t2m)

        def mlp(z, n):
            hid = F.gelu(F.linear(r(z, on["xn"]), sd[b + n + ".net.0.proj.weight"], sd[b + n + ".net.0.proj.bias"]), approximate="tanh")
            return F.linear(r(hid, on["hid"]), sd[b + n + ".net.2.weight"], sd[b + n + ".net.2.bias"])

        h = r(h + gi2 * mlp(xi2, "img_mlp"), on["res"])
        e = r(e + gt2 * mlp(xt2, "txt_mlp"), on["res"])
    emb = r(F.linear(st, sd["norm_out.linear.weight"], sd["norm_out.linear.bias"]), on["mod"])
    sc, sh = emb.chunk(2, 1)
    hn = r(O.layer_norm(h) * (1 + sc)[:, None] + sh[:, None], on["xn"])
    return r(F.linear(hn, sd["proj_out.weight"], sd["proj_out.bias"]), on["out"])


def main_gpu():
    """Same study at the REAL width and depth on the GPU (fp32 arithmetic, weights of the CUDA model upcast lazily),
    with the CUDA path's own result beside the emulation (does the rounding model explain the measured error?)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from depth_parity import LazyStateDict, build_model
    torch.backends.cuda.matmul.allow_tf32 = False
    tr, cfg = build_model(int(os.environ.get("L", "60")))
    sd = LazyStateDict(tr)
    dev, hh, s_txt = "cuda:0", 64, 512
    g = torch.Generator().manual_seed(11)
    lat = torch.randn(1, hh * hh, 64, generator=g).to(dev, torch.bfloat16)
    encb = torch.randn(1, s_txt, 3584, generator=g).to(dev, torch.bfloat16)
    t = torch.full((1,), 0.625, device=dev)
    shapes = [(1, hh, hh)]
    with torch.no_grad():
        ours = tr(hidden_states=lat, encoder_hidden_states=encb, timestep=t, img_shapes=[shapes], txt_seq_lens=[s_txt],
                  return_dict=False)[0].float()
        x, enc = lat.float(), encb.float()
        ref = O.transformer_forward(sd, cfg, x, enc, None, t, [shapes], [s_txt])
        print("CUDA path vs fp32 oracle                : %.3e" % O.rel_l2(ours, ref))
        cur = {p: True for p in POINTS}
        cur["res"] = False
        emu = forward(sd, cfg, x, enc, t, shapes, s_txt, cur)
        print("emulated roundings vs fp32 oracle       : %.3e   (CUDA path vs emulation %.3e)" % (O.rel_l2(emu, ref), O.rel_l2(ours, emu)))
        for p in POINTS:
            if p == "res":
                continue
            c = dict(cur)
            c[p] = False
            print("  current minus %-9s               : %.3e" % (p, O.rel_l2(forward(sd, cfg, x, enc, t, shapes, s_txt, c), ref)))
        for combo in (("mod", "temb"), ("xn", "hid"), ("qk_multi", "qkv", "p", "attn")):
            c = dict(cur)
            for p in combo:
                c[p] = False
            print("  current minus %-28s: %.3e" % ("+".join(combo), O.rel_l2(forward(sd, cfg, x, enc, t, shapes, s_txt, c), ref)))


def main():
    if torch.cuda.is_available() and "--gpu" in sys.argv:
        return main_gpu()
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = O.OracleConfig(num_layers=int(os.environ.get("L", "60")), num_attention_heads=4, joint_attention_dim=256)
    sd = {k: v.bfloat16().float() for k, v in O.init_state_dict(cfg, seed=5, bias_std=0.02, norm_weight_std=0.1).items()}
    hh = ww = 16
    s_txt = 64
    g = torch.Generator().manual_seed(11)
    x = torch.randn(1, hh * ww, 64, generator=g).bfloat16().float()
    enc = torch.randn(1, s_txt, 256, generator=g).bfloat16().float()
    t = torch.tensor([0.625])
    shapes = [(1, hh, ww)]
    with torch.no_grad():
        ref = O.transformer_forward(sd, cfg, x, enc, None, t, [shapes], [s_txt])
        none = forward(sd, cfg, x, enc, t, shapes, s_txt, {p: False for p in POINTS})
        print("self-check (all roundings off vs oracle): %.2e" % O.rel_l2(none, ref))
        cur = {p: True for p in POINTS}
        cur["res"] = False
        print("current CUDA path (fp32 streams)        : %.3e" % O.rel_l2(forward(sd, cfg, x, enc, t, shapes, s_txt, cur), ref))
        allon = {p: True for p in POINTS}
        print("round-1 path (bf16 streams)             : %.3e" % O.rel_l2(forward(sd, cfg, x, enc, t, shapes, s_txt, allon), ref))
        for p in POINTS:
            if p == "res":
                continue
            c = dict(cur)
            c[p] = False
            print("  current minus %-9s               : %.3e" % (p, O.rel_l2(forward(sd, cfg, x, enc, t, shapes, s_txt, c), ref)))
        for combo in (("mod", "temb"), ("mod", "temb", "qk_multi"), ("mod", "temb", "qk_multi", "out")):
            c = dict(cur)
            for p in combo:
                c[p] = False
            print("  current minus %-28s: %.3e" % ("+".join(combo), O.rel_l2(forward(sd, cfg, x, enc, t, shapes, s_txt, c), ref)))


if __name__ == "__main__":
    main()
