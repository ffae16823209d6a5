#!/usr/bin/env python
"""Per-operator error of ONE MMDiT block at full width (d=3072, 24x128 heads, 4096+512 tokens) on the GPU:
every intermediate the CUDA path leaves in its workspace vs the fp32 oracle's value of the same tensor.
Also cross-checks the fp32 oracle against itself in fp64 and with the math SDPA backend (is the checker exact?)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from depth_parity import LazyStateDict, build_model  # noqa: E402
from oracle import qwen_image_oracle as O  # noqa: E402


@torch.no_grad()
def main():
    dev = "cuda:0"
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    tr, ocfg = build_model(1)
    d, H, Dh = 3072, 24, 128
    s_img, s_txt, hh = 4096, 512, 64
    g = torch.Generator().manual_seed(11)
    lat = torch.randn(1, s_img, 64, generator=g).to(dev, torch.bfloat16)
    enc = torch.randn(1, s_txt, 3584, generator=g).to(dev, torch.bfloat16)
    mask = torch.ones(1, s_txt, dtype=torch.bool, device=dev)
    t = torch.full((1,), 0.625, device=dev)
    shapes = [[(1, hh, hh)]]
    cap = {}
    out = tr(hidden_states=lat, encoder_hidden_states=enc, encoder_hidden_states_mask=mask, timestep=t,
             img_shapes=shapes, txt_seq_lens=[s_txt], return_dict=False,
             _block_hook=lambda l, h, e: cap.update(h=h.clone(), e=e.clone()))[0]
    torch.cuda.synchronize()
    ws = next(iter(tr._ws.values()))
    res = {}

    def oracle(dtype, sdpa_math=False):
        sd = LazyStateDict(tr, dtype)
        x = lat.to(dtype)
        h = F.linear(x, sd["img_in.weight"], sd["img_in.bias"])
        e = F.linear(O.rms_norm(enc.to(dtype), sd["txt_norm.weight"]), sd["txt_in.weight"], sd["txt_in.bias"])
        temb = O.time_embed(sd, t.to(dtype), dtype)
        fi, ft = O.rope_freqs(ocfg, shapes[0], s_txt)
        fi, ft = fi.to(dev), ft.to(dev)
        if dtype == torch.float64:
            fi, ft = fi.to(torch.complex128), ft.to(torch.complex128)
        p = "transformer_blocks.0."
        im = F.linear(F.silu(temb), sd[p + "img_mod.1.weight"], sd[p + "img_mod.1.bias"])
        tm = F.linear(F.silu(temb), sd[p + "txt_mod.1.weight"], sd[p + "txt_mod.1.bias"])
        (i1, i2), (t1, t2) = im.chunk(2, -1), tm.chunk(2, -1)
        xi, gi1 = O._modulate(O.layer_norm(h), i1)
        xt, gt1 = O._modulate(O.layer_norm(e), t1)
        a = p + "attn."
        pr = lambda z, n: F.linear(z, sd[a + n + ".weight"], sd[a + n + ".bias"]).unflatten(-1, (H, Dh))

        def rope(z, fr):
            zc = torch.view_as_complex(z.reshape(*z.shape[:-1], -1, 2))
            return torch.view_as_real(zc * fr.unsqueeze(1)).flatten(3)

        iq, ik = rope(O.rms_norm(pr(xi, "to_q"), sd[a + "norm_q.weight"]), fi), rope(O.rms_norm(pr(xi, "to_k"), sd[a + "norm_k.weight"]), fi)
        tq, tk = rope(O.rms_norm(pr(xt, "add_q_proj"), sd[a + "norm_added_q.weight"]), ft), rope(O.rms_norm(pr(xt, "add_k_proj"), sd[a + "norm_added_k.weight"]), ft)
        iv, tv = pr(xi, "to_v"), pr(xt, "add_v_proj")
        q, k, v = torch.cat([tq, iq], 1), torch.cat([tk, ik], 1), torch.cat([tv, iv], 1)
        if sdpa_math or dtype == torch.float64:
            s = (q.transpose(1, 2) @ k.transpose(1, 2).transpose(-1, -2)) * Dh ** -0.5
            o = (torch.softmax(s, -1) @ v.transpose(1, 2)).transpose(1, 2)
        else:
            o = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)).transpose(1, 2)
        o = o.flatten(2)
        h1 = h + gi1 * F.linear(o[:, s_txt:], sd[a + "to_out.0.weight"], sd[a + "to_out.0.bias"])
        e1 = e + gt1 * F.linear(o[:, :s_txt], sd[a + "to_add_out.weight"], sd[a + "to_add_out.bias"])
        xi2, gi2 = O._modulate(O.layer_norm(h1), i2)
        hid = F.gelu(F.linear(xi2, sd[p + "img_mlp.net.0.proj.weight"], sd[p + "img_mlp.net.0.proj.bias"]), approximate="tanh")
        h2 = h1 + gi2 * F.linear(hid, sd[p + "img_mlp.net.2.weight"], sd[p + "img_mlp.net.2.bias"])
        return dict(q=q, k=k, v=v, attn=o, hid=hid, h=h2, h_after_attn=h1, h0=h, mod_img=im, temb=temb,
                    upd_attn=h1 - h, upd_mlp=h2 - h1)

    o32 = oracle(torch.float32)
    o32m = oracle(torch.float32, sdpa_math=True)
    o64 = oracle(torch.float64)
    for kname in ("q", "attn", "hid", "h", "upd_attn", "upd_mlp"):
        res[f"oracle32_vs_64.{kname}"] = O.rel_l2(o32[kname], o64[kname])
        res[f"oracle32math_vs_64.{kname}"] = O.rel_l2(o32m[kname], o64[kname])
    qkv = ws.qkv.float()
    ours = dict(q=qkv[:, :, :d].unflatten(-1, (H, Dh)), k=qkv[:, :, d:2 * d].unflatten(-1, (H, Dh)),
                v=qkv[:, :, 2 * d:].unflatten(-1, (H, Dh)), attn=ws.attn.float(), hid=ws.mlp_img.float(), h=cap["h"].float())
    for kname, val in ours.items():
        res[f"cuda_vs_64.{kname}"] = O.rel_l2(val, o64[kname])
    res["cuda_vs_64.h_minus_h0"] = O.rel_l2(ours["h"].double() - o64["h0"], o64["h"] - o64["h0"])
    res["rms"] = {k2: float(o64[k2].pow(2).mean().sqrt()) for k2 in ("h0", "h", "upd_attn", "upd_mlp", "attn", "hid", "mod_img", "temb")}
    print(json.dumps(res, indent=1))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "op_error_breakdown.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
