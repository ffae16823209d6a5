"""Property tests (hypothesis) for the integer / index work of the path — the part north_star wants bit-exact:
MSRoPE position ids and the patchify permutation through the C ABI against the oracle on random shapes, the
rank sharding arithmetic, and the scheduler table.  CPU only (host entry points of libqi_b200.so)."""
import pytest
import torch
from hypothesis import given, settings, strategies as st

from oracle import qwen_image_oracle as O


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as g
    g.build()


_shape = st.tuples(st.integers(1, 2), st.integers(1, 40), st.integers(1, 40))


@settings(max_examples=60, deadline=None)
@given(shapes=st.lists(_shape, min_size=1, max_size=3), txt=st.integers(0, 600))
def test_rope_position_ids_match_oracle_bitwise(shapes, txt):
    """(frame, h, w) ids of every image token (frame index = image index + frame, centred h/w ranges incl. odd
    sizes) and the text positions starting at max(h//2, w//2) over all images."""
    from vllm_omni_draft_b200 import ops
    ip, tp = ops.rope_position_ids(shapes, txt)
    oi, ot = O.rope_position_ids(shapes, txt)
    assert ip.dtype == torch.int32 and torch.equal(ip.long(), oi) and torch.equal(tp.long(), ot)
    assert ip.shape[0] == sum(f * h * w for f, h, w in shapes) and tp.shape[0] == txt
    # integer facts independent of the oracle: h/w ranges are centred, text continues after the largest half-extent
    k = 0
    for idx, (f, h, w) in enumerate(shapes):
        blk = ip[k:k + f * h * w].view(f, h, w, 3)
        assert int(blk[..., 0].min()) == idx and int(blk[..., 0].max()) == idx + f - 1
        assert int(blk[..., 1].min()) == -(h - h // 2) and int(blk[..., 1].max()) == h // 2 - 1
        assert int(blk[..., 2].min()) == -(w - w // 2) and int(blk[..., 2].max()) == w // 2 - 1
        k += f * h * w
    if txt:
        start = max(max(h // 2, w // 2) for _, h, w in shapes)
        assert tp[0].tolist() == [start] * 3 and tp[-1].tolist() == [start + txt - 1] * 3


@settings(max_examples=25, deadline=None)
@given(shapes=st.lists(_shape, min_size=1, max_size=2), txt=st.integers(1, 64))
def test_joint_rope_table_bitwise_and_unit_modulus(shapes, txt):
    from vllm_omni_draft_b200 import rope
    cfg = O.qwen_image_20b_config()
    tab = rope.joint_cos_sin(cfg.axes_dims_rope, cfg.rope_theta, shapes, txt)
    i, t = O.rope_freqs(cfg, shapes, txt)
    assert torch.equal(tab, torch.view_as_real(torch.cat([t, i], 0)))          # text rows first, bit-identical
    assert torch.allclose(tab.pow(2).sum(-1), torch.ones(tab.shape[:2]), atol=1e-6)


@settings(max_examples=40, deadline=None)
@given(c=st.sampled_from([4, 16]), hl=st.integers(1, 40), wl=st.integers(1, 40), b=st.integers(1, 3))
def test_pack_index_and_roundtrip(c, hl, wl, b):
    """The patchify permutation: C-ABI index table == oracle == product; unpack(pack(x)) == x."""
    from vllm_omni_draft_b200 import ops
    from vllm_omni_draft_b200.pipeline import pack_latents, unpack_latents
    hl, wl = 2 * hl, 2 * wl
    src = torch.arange(c * hl * wl, dtype=torch.int64).view(1, c, hl, wl)
    idx = ops.pack_latents_index(c, hl, wl)
    assert torch.equal(idx, O.pack_latents(src).reshape(-1)) and torch.equal(idx, pack_latents(src).reshape(-1))
    assert torch.equal(torch.sort(idx).values, torch.arange(c * hl * wl))       # a permutation
    x = torch.randn(b, c, 1, hl, wl)
    p = pack_latents(x.transpose(1, 2))
    assert p.shape == (b, (hl // 2) * (wl // 2), 4 * c)
    assert torch.equal(unpack_latents(p, hl, wl), x) and torch.equal(O.unpack_latents(p, hl, wl), x)


@settings(max_examples=200, deadline=None)
@given(n=st.integers(0, 20000), parts=st.integers(1, 8))
def test_shard_bounds_partition_the_rows(n, parts):
    from vllm_omni_draft_b200.launcher import shard_bounds
    from vllm_omni_draft_b200.peer import prefix, shard_counts
    counts = shard_counts(n, parts)
    assert sum(counts) == n and max(counts) - min(counts) <= 1 and counts == sorted(counts, reverse=True)
    pre = prefix(counts)
    bounds = [shard_bounds(n, parts, r) for r in range(parts)]
    assert bounds[0][0] == 0 and bounds[-1][1] == n
    assert all(bounds[r][1] == bounds[r + 1][0] for r in range(parts - 1))
    assert [b - a for a, b in bounds] == counts and pre[-1] == n


@settings(max_examples=100, deadline=None)
@given(world=st.sampled_from([1, 2, 3, 4, 6, 8]), do_cfg=st.booleans())
def test_parallel_layout_groups_are_consistent(world, do_cfg):
    from vllm_omni_draft_b200.launcher import ParallelLayout
    lays = [ParallelLayout.make(world, r, do_cfg) for r in range(world)]
    for lay in lays:
        assert lay.cfg_degree * lay.sp_degree == world and lay.cfg_degree == (2 if do_cfg and world % 2 == 0 else 1)
        assert lay.rank in lay.sp_ranks() and lay.rank in lay.pair_ranks()
        assert len(lay.sp_ranks()) == lay.sp_degree and len(lay.pair_ranks()) == lay.cfg_degree
        for r in lay.sp_ranks():          # same branch, distinct sequence shards
            assert lays[r].branch == lay.branch and lays[r].sp_ranks() == lay.sp_ranks()
        for r in lay.pair_ranks():        # same shard, the two CFG branches
            assert lays[r].sp_rank == lay.sp_rank
        assert sorted(lays[r].branch for r in lay.pair_ranks()) == list(range(lay.cfg_degree))
    with pytest.raises(ValueError):
        ParallelLayout.make(world, world)


@settings(max_examples=60, deadline=None)
@given(steps=st.integers(2, 100), seq=st.integers(16, 20000))
def test_schedule_table_matches_oracle_and_is_monotone(steps, seq):
    # (one step: sigmas = [1], upstream's terminal stretch divides 0 by 0 — NaN there, NaN here; not a use case)
    from vllm_omni_draft_b200.pipeline import FlowMatchEulerSchedule
    sch = FlowMatchEulerSchedule(steps, seq)
    ts, sig = O.flowmatch_sigmas(steps, seq)
    assert torch.equal(sch.sigmas, sig) and torch.equal(sch.timesteps, ts)
    assert sig[-1] == 0 and abs(float(sig[-2]) - 0.02) < 1e-5          # terminal stretch
    assert torch.all(sig[1:] <= sig[:-1]) and torch.all(sch.dts <= 0)
    assert abs(float(sch.dts.sum()) + float(sig[0])) < 1e-5            # the steps add up to -sigma_0
