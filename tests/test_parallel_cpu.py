"""world_size-2 gloo tests of the launcher's host-side logic (rank layout, shard bounds, Ulysses
head<->sequence exchange, CFG pair gather) — no GPU needed."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vllm_omni_draft_b200.launcher import ParallelLayout, UlyssesContext, shard_bounds


def test_layout_and_shards():
    lay = [ParallelLayout.make(8, r) for r in range(8)]
    assert [l.branch for l in lay] == [0, 0, 0, 0, 1, 1, 1, 1]
    assert [l.sp_rank for l in lay] == [0, 1, 2, 3, 0, 1, 2, 3]
    assert lay[5].sp_ranks() == [4, 5, 6, 7] and lay[5].pair_ranks() == [1, 5]
    l2 = ParallelLayout.make(2, 1)
    assert (l2.cfg_degree, l2.sp_degree, l2.branch) == (2, 1, 1)
    l1 = ParallelLayout.make(1, 0)
    assert (l1.cfg_degree, l1.sp_degree) == (1, 1)
    l4 = ParallelLayout.make(4, 2, do_cfg=False)
    assert (l4.cfg_degree, l4.sp_degree, l4.sp_rank) == (1, 4, 2)
    assert shard_bounds(4096, 4, 3) == (3072, 4096)
    # cfg-5's ragged length: 6889 = 1723 + 3 * 1722, shards tile the sequence without gaps
    b = [shard_bounds(6889, 4, i) for i in range(4)]
    assert b == [(0, 1723), (1723, 3445), (3445, 5167), (5167, 6889)]
    from vllm_omni_draft_b200.peer import prefix, shard_counts
    assert shard_counts(7401, 4) == [1851, 1850, 1850, 1850] and prefix([3, 2, 2]) == [0, 3, 5, 7]
    with pytest.raises(ValueError):
        ParallelLayout.make(2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, H, Dh, s_loc):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(123)
        full = torch.randn(world * s_loc, 3, H, Dh, generator=g)          # identical on every rank
        local = full[rank * s_loc:(rank + 1) * s_loc].reshape(s_loc, 3 * H * Dh).contiguous()
        ctx = UlyssesContext(dist.group.WORLD, world, rank)
        got = ctx.scatter_heads(local, H, Dh)                              # [S, 3, H/P, Dh]
        hs = H // world
        want = full[:, :, rank * hs:(rank + 1) * hs]
        assert torch.equal(got, want), "scatter_heads: wrong tokens/heads after all-to-all"
        # attention stand-in: o[s, h] = q[s, h] (identity) -> gather back to token-sharded, all heads
        o_full = got[:, 0].contiguous()                                    # [S, H/P, Dh]
        back = ctx.gather_heads(o_full)                                    # [P, S_loc, H/P*Dh]
        rebuilt = back.permute(1, 0, 2).reshape(s_loc, H * Dh)
        want_back = full[rank * s_loc:(rank + 1) * s_loc, 0].reshape(s_loc, H * Dh)
        assert torch.equal(rebuilt, want_back), "gather_heads: K-chunk order must be head-group major"
        # CFG pair gather: index 0 must be the cond rank's tensor
        pred = torch.full((1, 4, 64), float(rank))
        buf = torch.empty(2, 1, 4, 64)
        dist.all_gather_into_tensor(buf.view(-1), pred.reshape(-1))
        assert float(buf[0].mean()) == 0.0 and float(buf[1].mean()) == 1.0
    finally:
        dist.destroy_process_group()


def test_ulysses_exchange_world2_gloo():
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 4, 8, 6), nprocs=2, join=True)
