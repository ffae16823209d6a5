"""GPU parity of each C-ABI kernel against the oracle's op (same seeded inputs).
Tolerances: bf16 outputs -> rel-L2 <= 1e-2 (BASELINE.json north_star), usually far tighter."""
import math

import pytest
import torch

from oracle import qwen_image_oracle as O

pytestmark = pytest.mark.gpu


def _ops():
    from vllm_omni_draft_b200 import ops
    return ops


def _rand(shape, seed, scale=1.0, device="cuda"):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16).to(device)


def test_cfg_euler_step_matches_oracle(cuda_device):
    ops = _ops()
    B, S = 2, 4096 + 3
    c, u, x = _rand((B, S, 64), 1), _rand((B, S, 64), 2), _rand((B, S, 64), 3)
    sig = torch.tensor([0.7446114, 0.4266735])
    out = torch.empty_like(x)
    ops.cfg_euler_step(c, u, x, out, 4.0, dt=float(sig[1] - sig[0]))
    ref = O.scheduler_step(O.true_cfg_combine(c.cpu(), u.cpu(), 4.0), x.cpu(), sig[0], sig[1])
    diff = (out.cpu().float() - ref.float()).abs()
    # op-by-op bf16 rounding is reproduced; only fp32 summation order inside the norms may differ
    assert (diff > 0).float().mean() < 0.02, f"mismatch fraction {(diff > 0).float().mean()}"
    assert O.rel_l2(out.cpu(), ref) < 2e-3
    # no-CFG path and device dt
    dt_dev = torch.tensor([float(sig[1] - sig[0])], device="cuda")
    out2 = torch.empty_like(x)
    ops.cfg_euler_step(c, None, x, out2, 1.0, dt_dev=dt_dev)
    ref2 = O.scheduler_step(c.cpu(), x.cpu(), sig[0], sig[1])
    assert torch.equal(out2.cpu(), ref2)


@pytest.mark.parametrize("d", [128, 256, 3072, 3584])
def test_adaln_modulate(cuda_device, d):
    ops = _ops()
    B, S = 2, 333
    x = _rand((B * S, d), 4, 2.0)
    mod = _rand((B, 6 * d), 5, 0.3)
    out = torch.empty_like(x)
    ops.adaln_modulate(x, mod[:, 0:d], mod[:, d:2 * d], out, S)
    xf = x.cpu().float().view(B, S, d)
    ref = O.layer_norm(xf) * (1 + mod.cpu().float()[:, None, d:2 * d]) + mod.cpu().float()[:, None, 0:d]
    assert O.rel_l2(out.cpu().view(B, S, d), ref) < 4e-3


def test_rmsnorm_affine(cuda_device):
    ops = _ops()
    x, w = _rand((77, 3584), 6, 1.5), _rand((3584,), 7, 0.2) + 1
    out = torch.empty_like(x)
    ops.rmsnorm_affine(x, w, out)
    ref = O.rms_norm(x.cpu(), w.cpu())
    assert O.rel_l2(out.cpu(), ref) < 1e-3


@pytest.mark.parametrize("heads,hd,axes", [(4, 64, (8, 28, 28)), (24, 128, (16, 56, 56))])
def test_qknorm_rope(cuda_device, heads, hd, axes):
    ops = _ops()
    from vllm_omni_draft_b200 import rope
    B, s_txt, hh, ww = 2, 19, 6, 10
    s_img = hh * ww
    S = s_txt + s_img
    d = heads * hd
    qkv = _rand((B * S, 3 * d), 8, 1.3)
    ws = [_rand((hd,), 9 + i, 0.2) + 1 for i in range(4)]
    cfg = O.OracleConfig(attention_head_dim=hd, num_attention_heads=heads, axes_dims_rope=axes)
    img_f, txt_f = O.rope_freqs(cfg, [(1, hh, ww)], s_txt)
    freqs = rope.joint_cos_sin(axes, 10000.0, [(1, hh, ww)], s_txt).cuda()
    ref = qkv.cpu().clone().view(B, S, 3, heads, hd)
    wq_i, wk_i, wq_t, wk_t = [w.cpu() for w in ws]
    q, k = ref[:, :, 0], ref[:, :, 1]
    rq = torch.cat([O.apply_rope(O.rms_norm(q[:, :s_txt], wq_t), txt_f),
                    O.apply_rope(O.rms_norm(q[:, s_txt:], wq_i), img_f)], 1)
    rk = torch.cat([O.apply_rope(O.rms_norm(k[:, :s_txt], wk_t), txt_f),
                    O.apply_rope(O.rms_norm(k[:, s_txt:], wk_i), img_f)], 1)
    v_before = ref[:, :, 2].clone()
    ops.qknorm_rope(qkv, heads, hd, S, s_txt, ws[0], ws[1], ws[2], ws[3], freqs)
    got = qkv.cpu().view(B, S, 3, heads, hd)
    assert torch.equal(got[:, :, 2], v_before), "v must be untouched"
    assert O.rel_l2(got[:, :, 0], rq) < 2e-3
    assert O.rel_l2(got[:, :, 1], rk) < 2e-3


def test_small_linear_and_sinusoid(cuda_device):
    ops = _ops()
    x, w, b = _rand((3, 256), 20), _rand((512, 256), 21, 0.05), _rand((512,), 22, 0.1)
    out = torch.empty(3, 512, device="cuda", dtype=torch.bfloat16)
    ops.small_linear(x, w, b, out, silu_in=True)
    ref = torch.nn.functional.linear(torch.nn.functional.silu(x.cpu()).float(), w.cpu().float(), b.cpu().float())
    assert O.rel_l2(out.cpu(), ref) < 5e-3
    t = torch.tensor([1.0, 0.7446, 0.02], device="cuda")
    sin = torch.empty(3, 256, device="cuda", dtype=torch.bfloat16)
    ops.timestep_sinusoid(t, sin)
    assert O.rel_l2(sin.cpu(), O.timestep_sinusoid(t.cpu())) < 5e-3


GEMM_CASES = [
    # (B, M, N, K, epilogue)
    (1, 128, 128, 64, 0),
    (1, 128, 256, 256, 0),
    (1, 300, 768, 256, 0),
    (2, 200, 1024, 256, 1),
    (1, 4608, 3072, 3072, 2),
    (2, 257, 64, 1024, 0),
    (1, 4096, 9216, 3072, 0),
]


@pytest.mark.parametrize("B,M,N,K,epi", GEMM_CASES)
def test_gemm(cuda_device, B, M, N, K, epi):
    ops = _ops()
    a = _rand((B, M, K), 30, 1.0)
    w = _rand((N, K), 31, 0.05)
    bias = _rand((N,), 32, 0.5)
    out = torch.zeros(B, M, N, device="cuda", dtype=torch.bfloat16)
    res = _rand((B, M, N), 33) if epi == 2 else None
    gate = _rand((B, N), 34) if epi == 2 else None
    ops.gemm([ops.gemm_problem(a, w, bias, out, epi, res, gate)])
    torch.cuda.synchronize()
    y = torch.matmul(a.float(), w.float().t()) + bias.float()
    if epi == 1:
        y = torch.nn.functional.gelu(y, approximate="tanh")
    elif epi == 2:
        y = res.float() + gate.float()[:, None, :] * y
    err = O.rel_l2(out.cpu(), y.cpu())
    assert err < 5e-3, f"rel_l2={err}"


def test_gemm_two_problems_strided(cuda_device):
    """img + txt stream in one launch, writing into a joint [B, S, 3d] buffer (as the block does)."""
    ops = _ops()
    B, s_txt, s_img, d = 2, 40, 200, 256
    S = s_txt + s_img
    xi, xt = _rand((B, s_img, d), 40), _rand((B, s_txt, d), 41)
    wi, wt = _rand((3 * d, d), 42, 0.05), _rand((3 * d, d), 43, 0.05)
    bi, bt = _rand((3 * d,), 44), _rand((3 * d,), 45)
    joint = torch.zeros(B, S, 3 * d, device="cuda", dtype=torch.bfloat16)
    ops.gemm([ops.gemm_problem(xi, wi, bi, joint[:, s_txt:], 0),
              ops.gemm_problem(xt, wt, bt, joint[:, :s_txt], 0)])
    ref_i = xi.float() @ wi.float().t() + bi.float()
    ref_t = xt.float() @ wt.float().t() + bt.float()
    assert O.rel_l2(joint[:, s_txt:].cpu(), ref_i.cpu()) < 5e-3
    assert O.rel_l2(joint[:, :s_txt].cpu(), ref_t.cpu()) < 5e-3


ATT_CASES = [
    # (B, S, H, Dh)
    (1, 256, 2, 128),
    (1, 128, 1, 64),
    (2, 288, 4, 64),
    (1, 1000, 3, 128),
    (1, 4608, 24, 128),
]


@pytest.mark.parametrize("B,S,H,Dh", ATT_CASES)
def test_attention(cuda_device, B, S, H, Dh):
    ops = _ops()
    d = H * Dh
    qkv = _rand((B, S, 3 * d), 50, 1.0)
    q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
    out = torch.zeros(B, S, H, Dh, device="cuda", dtype=torch.bfloat16)
    ops.attention(q, k, v, out)
    torch.cuda.synchronize()
    ref = torch.nn.functional.scaled_dot_product_attention(
        q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2)).transpose(1, 2)
    err = O.rel_l2(out.cpu(), ref.cpu())
    assert err < 1e-2, f"rel_l2={err}"


@pytest.mark.parametrize("Dh", [128, 64])
@pytest.mark.parametrize("variant", [0, 3, 8, 11])
def test_attention_softmax_variants(cuda_device, variant, Dh):
    """Every shipped softmax form (MUFU only / 3 of 8 pairs on the FMA-pipe exp2 polynomial, P published whole
    or in two halves) matches fp32 SDPA, on a ragged sequence (partial
    last K/V tile, partial Q tile) with large-magnitude logits so the lazy rescale and the polynomial's range
    reduction are exercised."""
    ops = _ops()
    B, S, H = 2, 1000, 3
    d = H * Dh
    qkv = _rand((B, S, 3 * d), 51, 1.0)
    qkv[:, :, :d] *= 3.0                      # logits ~ N(0, 3*sqrt(128)*...) -> wide dynamic range
    q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
    out = torch.zeros(B, S, H, Dh, device="cuda", dtype=torch.bfloat16)
    ops.attention_set_variant(variant)
    try:
        ops.attention(q, k, v, out)
        torch.cuda.synchronize()
    finally:
        ops.attention_set_variant(-1)
    ref = torch.nn.functional.scaled_dot_product_attention(
        q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2)).transpose(1, 2)
    err = O.rel_l2(out.cpu(), ref.cpu())
    assert err < 6e-3, f"variant={variant} rel_l2={err}"


@pytest.mark.parametrize("Dh", [128, 64])
@pytest.mark.parametrize("S", [40, 64, 100, 128, 168, 300, 936, 1000, 1152])
def test_attention_ragged_lengths_moving_maximum(cuda_device, S, Dh):
    """Sequence ends everywhere relative to the 128-key K/V tile and the 256-row Q tile pair: shorter than one tile
    (40, 64, 100), on a boundary (128, 1152), inside a tile (168, 300, 936, 1000), an odd number of Q tiles (40, 300:
    the second tile of the pair has no rows to store); later keys score higher, so the running maximum keeps moving
    and the lazy O rescale has to wait for the P*V MMA still in flight.  Default form and the MUFU-only unsplit one."""
    ops = _ops()
    B, H = 2, 2
    d = H * Dh
    qkv = _rand((B, S, 3 * d), 52, 1.0)
    qkv[:, :, :d] *= 3.0
    qkv[:, S // 2:, d:2 * d] *= 2.0           # later keys score higher: the running maximum keeps moving
    q, k, v = [qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
    ref = torch.nn.functional.scaled_dot_product_attention(
        q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2)).transpose(1, 2)
    for variant in (-1, 0):
        out = torch.zeros(B, S, H, Dh, device="cuda", dtype=torch.bfloat16)
        ops.attention_set_variant(variant)
        try:
            ops.attention(q, k, v, out)
            torch.cuda.synchronize()
        finally:
            ops.attention_set_variant(-1)
        err = O.rel_l2(out.cpu(), ref.cpu())
        assert err < 6e-3, f"variant={variant} S={S} rel_l2={err}"


@pytest.mark.parametrize("bn", [128, 160, 192, 224, 256])
def test_gemm_tile_widths(cuda_device, bn):
    """Every tile width of the persistent GEMM (auto-selected by wave count in production) is exact."""
    ops = _ops()
    ops.gemm_set_tile_n(bn)
    try:
        for (B, M, N, K, epi) in [(1, 1152, 3072, 3072, 2), (2, 200, 1024, 256, 1), (1, 257, 64, 128, 0)]:
            a, w, bias = _rand((B, M, K), 60), _rand((N, K), 61, 0.05), _rand((N,), 62, 0.5)
            out = torch.zeros(B, M, N, device="cuda", dtype=torch.bfloat16)
            res = _rand((B, M, N), 63) if epi == 2 else None
            gate = _rand((B, N), 64) if epi == 2 else None
            ops.gemm([ops.gemm_problem(a, w, bias, out, epi, res, gate)])
            y = torch.matmul(a.float(), w.float().t()) + bias.float()
            if epi == 1:
                y = torch.nn.functional.gelu(y, approximate="tanh")
            elif epi == 2:
                y = res.float() + gate.float()[:, None, :] * y
            err = O.rel_l2(out.cpu(), y.cpu())
            assert err < 5e-3, f"bn={bn} case {(B, M, N, K, epi)} rel_l2={err}"
    finally:
        ops.gemm_set_tile_n(0)


def _peer_scatter(bases, prefix, me, chunk_stride=0):
    from vllm_omni_draft_b200._lib import PeerScatter
    d = PeerScatter()
    d.n_peers, d.my_index, d.chunk_stride = len(bases), me, chunk_stride
    for p, t in enumerate(bases):
        d.base[p] = t.data_ptr()
    for i, v in enumerate(prefix):
        d.row_prefix[i] = v
    return d


@pytest.mark.parametrize("P,B,counts", [(2, 1, [40, 40]), (4, 2, [37, 36, 36, 36])])
def test_ulysses_scatter_kernels_emulated_ranks(cuda_device, P, B, counts):
    """The fused exchange (K2 scatter -> attention scatter -> K-chunked out-proj operand) with the P
    ranks emulated on one GPU: every 'peer' buffer is a local tensor, so the layout arithmetic of both
    scatter kernels (ragged shards, batch, head ownership) is checked against the unsharded path."""
    ops = _ops()
    from vllm_omni_draft_b200 import rope
    H, Dh, axes = 4, 64, (8, 28, 28)
    d = H * Dh
    hs = H // P
    pre = [0]
    for c in counts:
        pre.append(pre[-1] + c)
    S = pre[-1]
    s_max = max(counts)
    s_txt_full = 2 * P + 1                       # ragged text split as well
    from vllm_omni_draft_b200.peer import prefix as _prefix, shard_counts
    txt_counts = shard_counts(s_txt_full, P)
    img_counts = [c - t for c, t in zip(counts, txt_counts)]
    s_img_full = sum(img_counts)
    tpre, ipre = _prefix(txt_counts), _prefix(img_counts)
    hh, ww = 1, s_img_full                       # one row of image tokens
    qkv_full = _rand((B, s_txt_full + s_img_full, 3 * d), 70, 1.2)          # joint order [txt; img]
    ws = [_rand((Dh,), 71 + i, 0.2) + 1 for i in range(4)]
    freqs_full = rope.joint_cos_sin(axes, 10000.0, [(1, hh, ww)], s_txt_full).cuda()
    # unsharded reference on the same kernels: K2 in place, then attention
    ref_qkv = qkv_full.clone()
    ops.qknorm_rope(ref_qkv.view(B * (s_txt_full + s_img_full), 3 * d), H, Dh, s_txt_full + s_img_full, s_txt_full,
                    ws[0], ws[1], ws[2], ws[3], freqs_full)
    rq, rk, rv = [ref_qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3)]
    ref_o = torch.zeros(B, s_txt_full + s_img_full, H, Dh, device="cuda", dtype=torch.bfloat16)
    ops.attention(rq, rk, rv, ref_o)
    # emulated ranks; exchange buffers are head-major: recv [B, 3, H/P, S, Dh], back [P*H/P, B, S_max, Dh]
    recv = [torch.zeros(B, 3, hs, S, Dh, device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    back = [torch.zeros(P * hs, B, s_max, Dh, device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    rows_of = []
    for r in range(P):
        rows = list(range(tpre[r], tpre[r + 1])) + [s_txt_full + i for i in range(ipre[r], ipre[r + 1])]
        rows_of.append(rows)
        local = qkv_full[:, rows].contiguous()                               # [B, S_loc, 3d]
        fl = freqs_full[rows].contiguous()
        ops.qknorm_rope_scatter(local.view(B * counts[r], 3 * d), H, Dh, counts[r], txt_counts[r],
                                ws[0], ws[1], ws[2], ws[3], fl, _peer_scatter(recv, pre, r))
    for r in range(P):
        full = recv[r]
        ops.attention_scatter(full[:, 0].transpose(1, 2), full[:, 1].transpose(1, 2), full[:, 2].transpose(1, 2),
                              _peer_scatter(back, pre, r, hs * B * s_max * Dh), Dh, s_max * Dh,
                              o_head_stride=B * s_max * Dh)
    torch.cuda.synchronize()
    for r in range(P):
        got = back[r][:, :, :counts[r]].permute(1, 2, 0, 3)                  # [B, S_loc, H (chunk = head), Dh]
        want = ref_o[:, rows_of[r]]
        err = O.rel_l2(got.cpu(), want.cpu())
        assert err < 4e-3, f"rank {r}: rel_l2={err}"
        # the scattered q/k/v themselves are bit-identical to the in-place kernel's output
        full_rows = [x for rr in rows_of for x in rr]
        for j, t in enumerate((rq, rk, rv)):
            assert torch.equal(recv[r][:, j].transpose(1, 2), t[:, full_rows][:, :, r * hs:(r + 1) * hs])
    # the out-projection reads `back` in place as a K-chunked A operand (one chunk of Dh columns per head)
    w = _rand((96, d), 77, 0.05)
    for r in range(P):
        out = torch.zeros(B, counts[r], 96, device="cuda", dtype=torch.bfloat16)
        ops.gemm([ops.gemm_problem(back[r][:, :, :counts[r]], w, None, out, 0, a_k_chunked=True)])
        a = back[r][:, :, :counts[r]].permute(1, 2, 0, 3).reshape(B, counts[r], d).float()
        assert O.rel_l2(out.cpu(), (a @ w.float().t()).cpu()) < 5e-3


@pytest.mark.parametrize("H,Dh,axes,bn,pair", [(4, 64, (8, 28, 28), 0, -1), (4, 64, (8, 28, 28), 192, 0),
                                               (4, 128, (16, 56, 56), 128, 1), (6, 128, (16, 56, 56), 256, 0),
                                               (6, 128, (16, 56, 56), 0, -1)])
def test_gemm_qkv_norm_rope_epilogue(cuda_device, H, Dh, axes, bn, pair):
    """K2 fused into K3: the QKV GEMM whose epilogue applies QK-RMSNorm + MSRoPE per head vector equals the
    plain QKV GEMM followed by qi_qknorm_rope (same rounding points; only the fp32 order of the 128-term sum
    of squares differs, so a few elements may move by one bf16 ulp), for both streams in one launch."""
    ops = _ops()
    from vllm_omni_draft_b200 import rope
    B, s_txt, hh, ww = 2, 37, 9, 31
    s_img = hh * ww
    S, d, K = s_txt + s_img, H * Dh, 256
    xi, xt = _rand((B, s_img, K), 80), _rand((B, s_txt, K), 81)
    wi, wt = _rand((3 * d, K), 82, 0.08), _rand((3 * d, K), 83, 0.08)
    bi, bt = _rand((3 * d,), 84, 0.3), _rand((3 * d,), 85, 0.3)
    nw = [_rand((Dh,), 86 + i, 0.2) + 1 for i in range(4)]
    freqs = rope.joint_cos_sin(axes, 10000.0, [(1, hh, ww)], s_txt).cuda()
    ref = torch.zeros(B, S, 3 * d, device="cuda", dtype=torch.bfloat16)
    ops.gemm([ops.gemm_problem(xi, wi, bi, ref[:, s_txt:]), ops.gemm_problem(xt, wt, bt, ref[:, :s_txt])])
    ops.qknorm_rope(ref.view(B * S, 3 * d), H, Dh, S, s_txt, nw[0], nw[1], nw[2], nw[3], freqs)
    out = torch.zeros_like(ref)
    ops.gemm_set_tile_n(bn)
    ops.gemm_set_cta_pair(pair)
    try:
        kw = dict(epilogue=ops.EPI_QKV_NORM_ROPE, freqs=freqs, heads=H, head_dim=Dh)
        ops.gemm([ops.gemm_problem(xi, wi, bi, out[:, s_txt:], qk_norm=(nw[0], nw[1]), rope_pos0=s_txt, **kw),
                  ops.gemm_problem(xt, wt, bt, out[:, :s_txt], qk_norm=(nw[2], nw[3]), rope_pos0=0, **kw)])
        torch.cuda.synchronize()
    finally:
        ops.gemm_set_tile_n(0)
        ops.gemm_set_cta_pair(-1)
    assert torch.equal(out[:, :, 2 * d:], ref[:, :, 2 * d:])            # v columns: plain bias epilogue
    err = O.rel_l2(out.float().cpu(), ref.float().cpu())
    same = (out == ref).float().mean().item()
    assert err < 2e-3 and same > 0.98, f"rel_l2={err} identical={same}"


@pytest.mark.parametrize("P,H,Dh,bn,pair,s_txt_full,s_img_full", [
    (2, 4, 64, 0, -1, 5, 71),          # schedule left to the heuristic
    (2, 4, 64, 128, 0, 5, 71),         # destination-ordered tiles: 1 column tile per (q|k|v, owner) group
    (2, 4, 64, 128, 1, 5, 71),
    (3, 6, 64, 128, 1, 7, 301),
    (4, 8, 128, 256, 1, 33, 700),      # the 8-GPU layout's schedule (pair kernel, 256-wide tiles), ragged rows
    (4, 8, 128, 128, 1, 33, 700),      # 2 column tiles per group
    (4, 8, 128, 128, 0, 33, 300),
    (4, 12, 128, 256, 1, 16, 520),     # 384 columns per group: not a multiple of 256 -> plain tile order
])
def test_gemm_qkv_epilogue_scatter_emulated_ranks(cuda_device, P, H, Dh, bn, pair, s_txt_full, s_img_full):
    """The QKV epilogue's Ulysses scatter (head vectors stored straight into the head owners' attention
    input buffers) against the in-place epilogue, P ranks emulated on one GPU, ragged shards, batch 2.
    With the scatter the column tiles are visited owner by owner (own heads last); every tile must still be
    produced exactly once, bit-identically."""
    ops = _ops()
    from vllm_omni_draft_b200 import rope
    from vllm_omni_draft_b200.peer import prefix as _prefix, shard_counts
    B, K = 2, 128
    axes = (8, 28, 28) if Dh == 64 else (16, 56, 56)
    d, hs = H * Dh, H // P
    tc, ic = shard_counts(s_txt_full, P), shard_counts(s_img_full, P)
    counts = [a + b for a, b in zip(tc, ic)]
    pre, tpre, ipre = _prefix(counts), _prefix(tc), _prefix(ic)
    S = pre[-1]
    x = _rand((B, S, K), 90)                     # joint order [txt; img]
    wi, wt = _rand((3 * d, K), 91, 0.1), _rand((3 * d, K), 92, 0.1)
    bi, bt = _rand((3 * d,), 93, 0.3), _rand((3 * d,), 94, 0.3)
    nw = [_rand((Dh,), 95 + i, 0.2) + 1 for i in range(4)]
    freqs_full = rope.joint_cos_sin(axes, 10000.0, [(1, 1, s_img_full)], s_txt_full).cuda()
    kw = dict(epilogue=ops.EPI_QKV_NORM_ROPE, heads=H, head_dim=Dh)
    ref = torch.zeros(B, S, 3 * d, device="cuda", dtype=torch.bfloat16)
    ops.gemm([ops.gemm_problem(x[:, s_txt_full:], wi, bi, ref[:, s_txt_full:], qk_norm=(nw[0], nw[1]),
                               rope_pos0=s_txt_full, freqs=freqs_full, **kw),
              ops.gemm_problem(x[:, :s_txt_full], wt, bt, ref[:, :s_txt_full], qk_norm=(nw[2], nw[3]),
                               rope_pos0=0, freqs=freqs_full, **kw)])
    recv = [torch.full((B, 3, hs, S, Dh), float("nan"), device="cuda", dtype=torch.bfloat16) for _ in range(P)]
    order = []
    ops.gemm_set_tile_n(bn)
    ops.gemm_set_cta_pair(pair)
    try:
        for r in range(P):
            rows = list(range(tpre[r], tpre[r + 1])) + [s_txt_full + i for i in range(ipre[r], ipre[r + 1])]
            order += rows
            xl = x[:, rows].contiguous()
            fl = freqs_full[rows].contiguous()
            sc = _peer_scatter(recv, pre, r)
            ops.gemm([ops.gemm_problem(xl[:, tc[r]:], wi, bi, None, qk_norm=(nw[0], nw[1]), rope_pos0=tc[r],
                                       freqs=fl, scatter=sc, **kw),
                      ops.gemm_problem(xl[:, :tc[r]], wt, bt, None, qk_norm=(nw[2], nw[3]), rope_pos0=0, freqs=fl,
                                       scatter=sc, **kw)])
        torch.cuda.synchronize()
    finally:
        ops.gemm_set_tile_n(0)
        ops.gemm_set_cta_pair(-1)
    want = ref[:, order].view(B, S, 3, H, Dh)
    for r in range(P):      # head-major destination [B, 3, H/P, S, Dh]; NaN anywhere = a tile never written
        assert torch.equal(recv[r].permute(0, 3, 1, 2, 4), want[:, :, :, r * hs:(r + 1) * hs])


def test_gemm_split_launch_matches_combined(cuda_device):
    """ops.gemm may issue one launch per problem when the autotuner measured that to be faster for these dims
    (a sequence-parallel rank's 128-row text stream wastes half of every 256-row pair tile in a combined launch).
    Same kernels, so the results must be bit-identical; the launch counter must see both launches."""
    ops = _ops()
    B, m1, m2, K, N = 1, 1024, 128, 256, 768
    a1, a2 = _rand((B, m1, K), 60), _rand((B, m2, K), 61)
    w1, w2 = _rand((N, K), 62, 0.05), _rand((N, K), 63, 0.05)
    b1, b2 = _rand((N,), 64, 0.5), _rand((N,), 65, 0.5)
    r1, r2 = _rand((B, m1, N), 66), _rand((B, m2, N), 67)
    gate = _rand((B, N), 68)

    def run():
        o1, o2 = torch.zeros_like(r1), torch.zeros_like(r2)
        probs = [ops.gemm_problem(a1, w1, b1, o1, ops.EPI_GATED_RESIDUAL, r1, gate),
                 ops.gemm_problem(a2, w2, b2, o2, ops.EPI_GATED_RESIDUAL, r2, gate)]
        n0 = ops.launch_count
        ops.gemm(probs)
        torch.cuda.synchronize()
        return o1, o2, ops.launch_count - n0, ops._gemm_key(probs), probs

    o1, o2, n, key, probs = run()
    assert n == 1 and key not in ops._split_keys
    ops._split_keys.add(key)
    try:
        s1, s2, n, _, _ = run()
    finally:
        ops._split_keys.discard(key)
    assert n == 2 and torch.equal(s1, o1) and torch.equal(s2, o2)
    y = r1.float() + gate.float()[:, None, :] * (a1.float() @ w1.float().t() + b1.float())
    assert O.rel_l2(o1.cpu(), y.cpu()) < 5e-3
    # the tuner runs on scratch outputs, registers single-problem schedules too, and keeps its books consistent
    ops.clear_gemm_schedules()
    chosen = ops.tune_gemm(probs, allow_split=True)
    assert chosen is not None and ops.tune_gemm(probs, allow_split=True) is None
    assert ops._gemm_key(probs[:1]) in ops._tuned_keys and ops._gemm_key(probs[1:]) in ops._tuned_keys
    t1, t2, n, _, _ = run()
    assert n == (2 if key in ops._split_keys else 1)
    assert O.rel_l2(t1.cpu(), y.cpu()) < 5e-3
    ops.clear_gemm_schedules()
    assert not ops._split_keys and not ops._tuned_keys


def test_peer_barrier_single_rank(cuda_device):
    """n_peers = 1: the rank signals itself; the epoch advances and the timeout flag stays clear."""
    ops = _ops()
    sig = torch.zeros(8, dtype=torch.int32, device="cuda")
    table = torch.tensor([sig.data_ptr()], dtype=torch.int64, device="cuda")
    state = torch.zeros(2, dtype=torch.int32, device="cuda")
    for _ in range(3):
        ops.peer_barrier(table, 1, 0, state)
    torch.cuda.synchronize()
    assert state.tolist() == [3, 0] and int(sig[0]) == 3


@pytest.mark.parametrize("bn", [128, 160, 192, 224, 256])
def test_gemm_cta_pair(cuda_device, bn):
    """tcgen05 cta_group::2 kernel (256 x BN tile per 2-CTA cluster): ragged M, batch, both fused
    epilogues, two problems in one launch and the K-chunked A operand."""
    ops = _ops()
    ops.gemm_set_tile_n(bn)
    ops.gemm_set_cta_pair(1)
    try:
        for (B, M, N, K, epi) in [(1, 4608, 3072, 3072, 2), (2, 300, 1024, 256, 1), (1, 257, 64, 128, 0),
                                  (1, 1152, 9216, 3072, 0)]:
            a, w, bias = _rand((B, M, K), 80), _rand((N, K), 81, 0.05), _rand((N,), 82, 0.5)
            out = torch.zeros(B, M, N, device="cuda", dtype=torch.bfloat16)
            res = _rand((B, M, N), 83) if epi == 2 else None
            gate = _rand((B, N), 84) if epi == 2 else None
            ops.gemm([ops.gemm_problem(a, w, bias, out, epi, res, gate)])
            y = torch.matmul(a.float(), w.float().t()) + bias.float()
            if epi == 1:
                y = torch.nn.functional.gelu(y, approximate="tanh")
            elif epi == 2:
                y = res.float() + gate.float()[:, None, :] * y
            err = O.rel_l2(out.cpu(), y.cpu())
            assert err < 5e-3, f"pair bn={bn} case {(B, M, N, K, epi)} rel_l2={err}"
        # two problems + K-chunked A ([chunks, B, M, Kc])
        P_, Bb, M1, M2, kc, N = 4, 2, 200, 56, 64, 512
        a1, a2 = _rand((P_, Bb, M1, kc), 85), _rand((P_, Bb, M2, kc), 86)
        w1, w2 = _rand((N, P_ * kc), 87, 0.05), _rand((N, P_ * kc), 88, 0.05)
        o1 = torch.zeros(Bb, M1, N, device="cuda", dtype=torch.bfloat16)
        o2 = torch.zeros(Bb, M2, N, device="cuda", dtype=torch.bfloat16)
        ops.gemm([ops.gemm_problem(a1, w1, None, o1, 0, a_k_chunked=True),
                  ops.gemm_problem(a2, w2, None, o2, 0, a_k_chunked=True)])
        for a, w, o in ((a1, w1, o1), (a2, w2, o2)):
            af = a.permute(1, 2, 0, 3).reshape(Bb, a.shape[2], P_ * kc).float()
            assert O.rel_l2(o.cpu(), (af @ w.float().t()).cpu()) < 5e-3
    finally:
        ops.gemm_set_tile_n(0)
        ops.gemm_set_cta_pair(-1)


# ------------------------------------------------------------------ fp32 residual streams (c_fp32 / x_fp32)
@pytest.mark.parametrize("bn,pair", [(0, -1), (128, 0), (160, 1), (192, 0), (224, 1), (256, 1)])
@pytest.mark.parametrize("epi", [0, 2])
def test_gemm_fp32_output_and_residual(cuda_device, bn, pair, epi):
    """C (and R) in fp32: the residual-stream form of the bias and gated-residual epilogues, every tile width,
    ragged M and N, two problems in one launch, in place (R aliases C)."""
    ops = _ops()
    B, K = 2, 512
    shapes = [(333, 1096), (77, 1096)]
    probs, refs, outs, keep = [], [], [], []          # `keep`: a GemmProblem holds raw pointers, not references
    for i, (M, N) in enumerate(shapes):
        a, w, bias = _rand((B, M, K), 60 + i), _rand((N, K), 62 + i, 0.05), _rand((N,), 64 + i, 0.5)
        keep += [a, w, bias]
        y = torch.matmul(a.float(), w.float().t()) + bias.float()
        if epi == 2:
            g = torch.Generator().manual_seed(70 + i)
            res = torch.randn(B, M, N, generator=g).cuda()                  # fp32 stream, not bf16-representable
            gate = _rand((B, N), 66 + i)
            y = res + gate.float()[:, None, :] * y
            out = res.clone()
            keep.append(gate)
            probs.append(ops.gemm_problem(a, w, bias, out, epi, out, gate))
        else:
            out = torch.zeros(B, M, N, device="cuda", dtype=torch.float32)
            probs.append(ops.gemm_problem(a, w, bias, out, epi))
        refs.append(y)
        outs.append(out)
    try:
        ops.gemm_set_tile_n(bn)
        ops.gemm_set_cta_pair(pair)
        ops.gemm(probs)
        torch.cuda.synchronize()
    finally:
        ops.gemm_set_tile_n(0)
        ops.gemm_set_cta_pair(-1)
    for out, y in zip(outs, refs):
        # fp32 in, fp32 out: only the bf16 operands' products (exact in fp32) and summation order differ
        assert O.rel_l2(out, y) < 1e-5


@pytest.mark.parametrize("mod_f32", [False, True])
@pytest.mark.parametrize("d", [128, 256, 3072])
def test_adaln_modulate_fp32_rows(cuda_device, d, mod_f32):
    ops = _ops()
    B, S1, S2 = 2, 333, 41
    g = torch.Generator().manual_seed(9)
    x1, x2 = (torch.randn(B * S1, d, generator=g) * 2).cuda(), (torch.randn(B * S2, d, generator=g) * 2).cuda()
    mod = _rand((B, 6 * d), 5, 0.3)
    if mod_f32:
        mod = (torch.randn(B, 6 * d, generator=g) * 0.3).cuda()
    o1 = torch.empty(B * S1, d, device="cuda", dtype=torch.bfloat16)
    o2 = torch.empty(B * S2, d, device="cuda", dtype=torch.bfloat16)
    ops.adaln_modulate_pair([(x1, mod[:, 0:d], mod[:, d:2 * d], o1, S1), (x2, mod[:, 3 * d:4 * d], mod[:, 4 * d:5 * d], o2, S2)])
    m = mod.float()
    for x, o, S, c0 in ((x1, o1, S1, 0), (x2, o2, S2, 3 * d)):
        ref = O.layer_norm(x.view(B, S, d)) * (1 + m[:, None, c0 + d:c0 + 2 * d]) + m[:, None, c0:c0 + d]
        assert O.rel_l2(o.view(B, S, d), ref) < 3e-3            # one bf16 rounding of the result


def test_time_embed_and_modulation_fp32(cuda_device):
    """qi_timestep_sinusoid_f32 + qi_small_linear_f32: the fp32 timestep-embedding / modulation chain vs PyTorch fp32."""
    ops = _ops()
    M, d, n = 5, 256, 1544
    t = torch.tensor([1.0, 0.7446, 0.5, 0.1, 0.02], device="cuda")
    w1, b1, w2, b2, wm, bm = (_rand((d, 256), 90, 0.05), _rand((d,), 91, 0.1), _rand((d, d), 92, 0.05), _rand((d,), 93, 0.1),
                              _rand((n, d), 94, 0.05), _rand((n,), 95, 0.1))
    sin = torch.empty(M, 256, device="cuda")
    t1, temb, mod = torch.empty(M, d, device="cuda"), torch.empty(M, d, device="cuda"), torch.empty(M, n, device="cuda")
    ops.timestep_sinusoid(t, sin)
    ops.small_linear(sin, w1, b1, t1, silu_out=True)
    ops.small_linear(t1, w2, b2, temb)
    ops.small_linear(temb, wm, bm, mod, silu_in=True)
    F = torch.nn.functional
    ref_sin = O.timestep_sinusoid(t)
    ref_temb = F.linear(F.silu(F.linear(ref_sin, w1.float(), b1.float())), w2.float(), b2.float())
    ref_mod = F.linear(F.silu(ref_temb), wm.float(), bm.float())
    # cos/sin of arguments up to 1000: one ulp of the frequency moves the argument by 6e-5
    assert O.rel_l2(sin, ref_sin) < 2e-4
    assert O.rel_l2(temb, ref_temb) < 2e-4
    assert O.rel_l2(mod, ref_mod) < 2e-4
    # given the same sinusoid the two Linears themselves are exact to fp32 summation order
    ref2 = F.linear(F.silu(F.linear(F.silu(F.linear(sin, w1.float(), b1.float())), w2.float(), b2.float())), wm.float(), bm.float())
    assert O.rel_l2(mod, ref2) < 5e-6


# ------------------------------------------------------------------ torch.ops.qi (dispatcher-visible C ABI)
def test_torch_ops_inference_mode_and_cuda_graph(cuda_device):
    """the C ABI registered as torch.library ops — callable under inference_mode and capturable."""
    import vllm_omni_draft_b200.torch_ops as T
    Q = torch.ops.qi
    assert set(T.OP_NAMES) >= {"gemm_bf16", "attention_fwd", "adaln_modulate", "qknorm_rope", "cfg_euler_step"}
    B, S, H, Dh, d = 1, 256, 2, 128, 256
    with torch.inference_mode():
        a, w, bias = _rand((B, S, d), 80), _rand((3 * d, d), 81, 0.05), _rand((3 * d,), 82)
        qkv = torch.zeros(B, S, 3 * d, device="cuda", dtype=torch.bfloat16)
        o = torch.zeros(B, S, H, Dh, device="cuda", dtype=torch.bfloat16)
        h = torch.randn(B, S, d, device="cuda")
        gate, wo, bo = _rand((B, d), 83), _rand((d, d), 84, 0.05), _rand((d,), 85)
        xn = torch.empty(B * S, d, device="cuda", dtype=torch.bfloat16)
        mod = _rand((B, 2 * d), 86, 0.3)
        c, u, x = _rand((B, S, 64), 87), _rand((B, S, 64), 88), _rand((B, S, 64), 89)
        nxt = torch.empty_like(x)

        def body():
            Q.adaln_modulate([h.view(B * S, d)], [mod[:, :d]], [mod[:, d:]], [xn], [S])
            Q.gemm_bf16([xn.view(B, S, d)], [w], [bias], [qkv], 0, [None], [None])
            q, k, v = (qkv[:, :, i * d:(i + 1) * d].unflatten(-1, (H, Dh)) for i in range(3))
            Q.attention_fwd(q, k, v, o)
            Q.gemm_bf16([o.view(B, S, d)], [wo], [bo], [h], 2, [h], [gate])
            Q.cfg_euler_step(c, u, x, nxt, 4.0, None, -0.05)

        h0 = h.clone()
        body()
        torch.cuda.synchronize()
        eager_h, eager_n = h.clone(), nxt.clone()
        h.copy_(h0)
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            body()
        torch.cuda.current_stream().wait_stream(side)
        h.copy_(h0)
        with torch.cuda.graph(g):
            body()
        h.copy_(h0)
        nxt.zero_()
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(h, eager_h) and torch.equal(nxt, eager_n)
        # against fp32 PyTorch
        xr = O.layer_norm(h0) * (1 + mod.float()[:, None, d:]) + mod.float()[:, None, :d]
        qkv_r = xr.bfloat16().float() @ w.float().t() + bias.float()
        qr, kr, vr = (qkv_r[:, :, i * d:(i + 1) * d].bfloat16().float().unflatten(-1, (H, Dh)).transpose(1, 2) for i in range(3))
        orf = torch.nn.functional.scaled_dot_product_attention(qr, kr, vr).transpose(1, 2).flatten(2)
        hr = h0 + gate.float()[:, None] * (orf.bfloat16().float() @ wo.float().t() + bo.float())
        assert O.rel_l2(eager_h, hr) < 3e-3
    with pytest.raises((NotImplementedError, RuntimeError)):
        Q.attention_fwd(torch.zeros(1, 8, 1, 64), torch.zeros(1, 8, 1, 64), torch.zeros(1, 8, 1, 64),
                        torch.zeros(1, 8, 1, 64))          # CPU tensors: no such kernel, no fallback


def test_cfg_euler_step_f32_matches_oracle(cuda_device):
    """K5 in fp32 (the pipeline's form): fp32 predictions and latents, in place, + the bf16 copy for the next forward."""
    ops = _ops()
    B, S = 2, 4096 + 3
    g = torch.Generator().manual_seed(4)
    c, u, x = (torch.randn(B, S, 64, generator=g).cuda() for _ in range(3))
    sig = torch.tensor([0.7446114, 0.4266735])
    ref = O.scheduler_step(O.true_cfg_combine(c, u, 4.0), x, sig[0].cuda(), sig[1].cuda())
    xb = torch.empty(B, S, 64, device="cuda", dtype=torch.bfloat16)
    x1 = x.clone()
    ops.cfg_euler_step_f32(c, u, x1, x1, xb, 4.0, dt=float(sig[1] - sig[0]))       # in place
    assert O.rel_l2(x1, ref) < 1e-6
    assert torch.equal(xb, x1.bfloat16())
    out = torch.empty_like(x)
    ops.cfg_euler_step_f32(c, None, x, out, None, 1.0, dt_dev=(sig[1] - sig[0]).reshape(1).cuda())
    assert O.rel_l2(out, O.scheduler_step(c, x, sig[0].cuda(), sig[1].cuda())) < 1e-6
