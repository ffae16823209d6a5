"""CFG-parallel / Ulysses sequence-parallel launcher (one process per GPU, torch.distributed + NCCL).

The reference has no parallelism of any kind; this is the partitioning
BASELINE.json's north_star asks for:

* CFG-parallel (2 branches): ranks of branch 0 run the cond forward, branch 1 the uncond forward;
  ONE collective per step — an all-gather of the local noise-prediction shard inside each
  (cond, uncond) pair — after which both ranks run the fused CFG+Euler kernel redundantly on
  identical inputs, so the latents never need a broadcast.
* Ulysses SP (degree P inside a branch): the joint [txt; img] token sequence is sharded over P
  ranks; every op except attention is token-local.  Around attention the head<->sequence all-to-all
  turns [S/P tokens, H heads] into [S tokens, H/P heads] and back (NCCL all_to_all_single over
  NVLink).  Key order after the exchange is rank-major, which is harmless: attention without a mask
  is invariant to key order and RoPE has already been applied.

Rank layout for world = 2 * P:  branch = rank // P,  sp_rank = rank % P.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def init_distributed(n_gpus: int) -> Tuple[int, int, int]:
    """Reads RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* (torchrun); returns (world, rank, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    if n_gpus != world and world > 1:
        raise ValueError(f"--gpus {n_gpus} does not match WORLD_SIZE {world}")
    return world, rank, local_rank


@dataclass
class ParallelLayout:
    world: int
    rank: int
    cfg_degree: int       # 1 or 2
    sp_degree: int
    branch: int           # 0 cond, 1 uncond
    sp_rank: int

    @staticmethod
    def make(world: int, rank: int, do_cfg: bool = True) -> "ParallelLayout":
        if world < 1 or rank < 0 or rank >= world:
            raise ValueError("bad world/rank")
        cfg_degree = 2 if (do_cfg and world >= 2 and world % 2 == 0) else 1
        sp = world // cfg_degree
        return ParallelLayout(world, rank, cfg_degree, sp, rank // sp if cfg_degree == 2 else 0, rank % sp)

    def sp_ranks(self) -> List[int]:
        base = self.branch * self.sp_degree if self.cfg_degree == 2 else 0
        return list(range(base, base + self.sp_degree))

    def pair_ranks(self) -> List[int]:
        return [self.sp_rank, self.sp_degree + self.sp_rank] if self.cfg_degree == 2 else [self.rank]


def shard_bounds(n: int, parts: int, idx: int) -> Tuple[int, int]:
    """Rows [lo, hi) of rank idx when n rows are split over `parts` ranks; ragged splits give the
    first n % parts ranks one extra row (cfg-5: 6889 image tokens over 4 ranks)."""
    from .peer import prefix, shard_counts
    pre = prefix(shard_counts(n, parts))
    return pre[idx], pre[idx + 1]


# ------------------------------------------------------------------ Ulysses exchanges (backend-agnostic)
class UlyssesContext:
    """head<->sequence all-to-all around attention for one sequence-parallel group."""

    def __init__(self, group, size: int, rank: int):
        self.group, self.size, self.rank = group, size, rank
        self.peer = None      # peer.PeerExchange: the fused peer-memory exchange replaces both all-to-alls

    def scatter_heads(self, qkv_local: torch.Tensor, heads: int, head_dim: int,
                      send: Optional[torch.Tensor] = None, recv: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[S_loc, 3*H*Dh] (token-sharded, all heads) -> [P*S_loc, 3, H/P, Dh] (all tokens, rank-major
        order, local heads)."""
        P = self.size
        s_loc = qkv_local.shape[0]
        hs = heads // P
        src = qkv_local.view(s_loc, 3, P, hs, head_dim).permute(2, 0, 1, 3, 4)
        if send is None:
            send = torch.empty(P, s_loc, 3, hs, head_dim, device=qkv_local.device, dtype=qkv_local.dtype)
        send.copy_(src)
        if recv is None:
            recv = torch.empty_like(send)
        dist.all_to_all_single(recv.view(-1), send.view(-1), group=self.group)
        return recv.view(P * s_loc, 3, hs, head_dim)

    def gather_heads(self, o_full: torch.Tensor, recv: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[P*S_loc, H/P, Dh] (all tokens, local heads) -> [P, S_loc, H/P*Dh]: chunk p holds head group
        p of the local tokens (the out-projection GEMM reads it as a K-chunked A operand)."""
        P = self.size
        s_loc = o_full.shape[0] // P
        send = o_full.reshape(P, s_loc, -1)
        if recv is None:
            recv = torch.empty_like(send)
        dist.all_to_all_single(recv.view(-1), send.view(-1), group=self.group)
        return recv


# ------------------------------------------------------------------ denoiser
class ParallelDenoiser:
    """Owns one rank's model replica and denoising session; `step(i)` advances the (local shard of
    the) latents by one scheduler step on every rank."""

    def __init__(self, cfg, device, world: int = 1, rank: int = 0, seed: int = 0, use_cuda_graph: bool = True,
                 state_dict=None, cfg_parallel: Optional[bool] = None, exchange: Optional[str] = None):
        """cfg_parallel: split cond/uncond over two rank groups (default: yes when world is even; the env
        QI_CFG_PARALLEL=0 gives pure Ulysses over all ranks).  exchange: "peer" (fused peer-memory
        scatter kernels, default) or "nccl" (all_to_all_single), env QI_SP_EXCHANGE."""
        from .pipeline import B200QwenImagePipeline
        from .transformer import B200QwenImageTransformer2DModel
        self.cfg, self.device, self.world, self.rank = cfg, device, world, rank
        if cfg_parallel is None:
            cfg_parallel = os.environ.get("QI_CFG_PARALLEL", "1") != "0"
        self.exchange = (exchange or os.environ.get("QI_SP_EXCHANGE", "peer")).lower()
        if self.exchange not in ("peer", "nccl"):
            raise ValueError("exchange must be 'peer' or 'nccl'")
        self.layout = ParallelLayout.make(world, rank, do_cfg=cfg_parallel)
        self.peer = None
        self.tr = B200QwenImageTransformer2DModel(cfg, device=device)
        if state_dict is not None:
            self.tr.load_diffusers_state_dict(state_dict)
        else:
            self.tr.init_random_(seed)            # same seed on every rank -> replicated weights
        self.pipe = B200QwenImagePipeline(self.tr)
        if world > 1:
            self.pipe.parallel = self        # pipe(...) on every rank = one sharded denoise (see `denoise`)
        self.use_cuda_graph = use_cuda_graph
        # peer flag-barrier timeout: a rank that died or diverged must surface as an error, not as a hang; host
        # skew at setup (per-rank autotuning, graph capture) is absorbed by explicit barriers, not by this
        self.peer_timeout_ms = int(os.environ.get("QI_PEER_TIMEOUT_MS", "10000"))
        self._setup_key = None
        self.graph = None
        self.sp_group = self.pair_group = None
        if world > 1:
            lay = self.layout
            # every rank must create every group, in the same order
            for b in range(lay.cfg_degree):
                ranks = list(range(b * lay.sp_degree, (b + 1) * lay.sp_degree))
                g = dist.new_group(ranks)
                if b == lay.branch:
                    self.sp_group = g
            if lay.cfg_degree == 2:
                for s in range(lay.sp_degree):
                    g = dist.new_group([s, lay.sp_degree + s])
                    if s == lay.sp_rank:
                        self.pair_group = g
            if lay.sp_degree > 1:
                self.tr.set_sequence_parallel(UlyssesContext(self.sp_group, lay.sp_degree, lay.sp_rank))
        self.sess = None
        self.launches_per_step = None

    # -------------------------------------------------------------- setup / IO
    def setup(self, cond_h, uncond_h, lat_h, height, width, num_inference_steps, true_cfg_scale=4.0,
              ref_h=None, ref_hw=None, sigmas=None):
        """ref_h [B, S_ref, 64] + ref_hw = (h, w) token grid: edit mode — the packed reference-image
        latents ride behind the noise tokens; under sequence parallelism noise and reference tokens are sharded
        separately, so every rank updates an equal share of the latents."""
        from . import ops
        from .pipeline import DenoiseSession, latent_hw
        lay, dev, dt = self.layout, self.device, self.tr.dtype
        # a previous setup's graph and session hold pointers into the previous peer buffers: drop them first
        self.graph = None
        if self.sess is not None:
            self.sess.close()
            self.sess = None
        self._setup_key = None
        Hl, Wl = latent_hw(height, width)
        self.s_img, self.s_txt = lat_h.shape[1], cond_h.shape[1]
        self.i0, self.i1 = shard_bounds(self.s_img, lay.sp_degree, lay.sp_rank)
        self.t0, self.t1 = shard_bounds(self.s_txt, lay.sp_degree, lay.sp_rank)
        s_ref = 0 if ref_h is None else ref_h.shape[1]
        if s_ref and (ref_hw is None or ref_hw[0] * ref_hw[1] != s_ref):
            raise ValueError("reference-image tokens do not match their (h, w) token grid")
        j0, j1 = shard_bounds(s_ref, lay.sp_degree, lay.sp_rank)
        self.j0, self.j1 = j0, j1
        ref = ref_h.to(dev, dt)[:, j0:j1].contiguous() if s_ref else None
        ref_kw = dict(ref_tokens=ref, ref_hw=ref_hw) if s_ref else {}
        if sigmas is not None:          # custom schedule (upstream's `sigmas=`): same list on every rank
            ref_kw["sigmas"] = [float(v) for v in sigmas]
            num_inference_steps = len(ref_kw["sigmas"])
        self.scale = float(true_cfg_scale)
        if lay.cfg_degree == 1:
            cond, uncond = cond_h.to(dev, dt), (uncond_h.to(dev, dt) if uncond_h is not None else None)
        else:
            mine = cond_h if lay.branch == 0 else uncond_h
            cond, uncond = mine.to(dev, dt), None
        cond = cond[:, self.t0:self.t1].contiguous()
        if uncond is not None:
            uncond = uncond[:, self.t0:self.t1].contiguous()
        lat = lat_h.to(dev)[:, self.i0:self.i1].contiguous()       # dtype kept: it seeds the fp32 master copy
        # stacked (cond, uncond) forward whenever both branches run on this rank; the NCCL exchange
        # is batch-1 only, so it keeps two forwards
        batched = not (lay.sp_degree > 1 and self.exchange == "nccl")
        if lay.sp_degree > 1:
            self.tr.set_sequence_shard((1, Hl // 2, Wl // 2), self.s_img, self.s_txt, (self.i0, self.i1),
                                       (self.t0, self.t1), ref_shape=(1, ref_hw[0], ref_hw[1]) if s_ref else None,
                                       ref_range=(j0, j1))
            if self.exchange == "peer":
                from .peer import PeerExchange, shard_counts
                rows = [a + b + c for a, b, c in zip(shard_counts(self.s_txt, lay.sp_degree),
                                                     shard_counts(self.s_img, lay.sp_degree),
                                                     shard_counts(s_ref, lay.sp_degree))]
                if self.peer is not None:
                    self.peer.close()
                self.peer = PeerExchange(self.sp_group, lay.sp_degree, lay.sp_rank, dev,
                                         self.cfg.num_attention_heads, self.cfg.attention_head_dim,
                                         lat.shape[0] * (2 if (batched and lay.cfg_degree == 1 and uncond is not None) else 1),
                                         rows, timeout_ms=self.peer_timeout_ms)
                self.tr._sp.peer = self.peer
            elif (self.s_img % lay.sp_degree) or (self.s_txt % lay.sp_degree) or (s_ref % lay.sp_degree):
                raise NotImplementedError("ragged shards need the peer-memory exchange (exchange='peer')")
        if self.world == 1:
            n0 = ops.launch_count
            self.sess = DenoiseSession(self.pipe, cond, uncond, lat, Hl, Wl, num_inference_steps, self.scale,
                                       use_cuda_graph=self.use_cuda_graph, **ref_kw)
            if self.use_cuda_graph:
                # launches recorded while capturing one step (+2 refresh copies are torch memcpys)
                self.launches_per_step = (ops.launch_count - n0 - 4) // 2
        else:
            self.sess = DenoiseSession(self.pipe, cond, uncond, lat, Hl, Wl, num_inference_steps, self.scale,
                                       use_cuda_graph=False, image_seq_len=self.s_img, batched_cfg=batched, **ref_kw)
            self.pair_buf = torch.empty(2, *lat.shape[:-1], 64, device=dev,
                                        dtype=torch.float32 if self.sess.precise else dt)
            if self.use_cuda_graph:
                # One graph per rank = forward + pair all-gather + K5 (+ the Ulysses all-to-alls inside the
                # forward).  NCCL communicators are created by an eager warm-up step first; capture is
                # thread-local so the NCCL watchdog thread cannot invalidate it.
                tr, s = self.tr, self.sess
                self.mod_cur = torch.empty(1, tr.n_mod, device=dev, dtype=s.mod_table.dtype)
                self.dt_cur = torch.empty(1, device=dev, dtype=torch.float32)
                self.mod_cur.copy_(s.mod_table[0:1])
                self.dt_cur.copy_(s.dts[0:1])
                keep = s.latents_master.clone()
                # every rank has finished its own autotuning before anyone enters the first flag barrier
                torch.cuda.synchronize()
                dist.barrier()
                self._step_body(self.mod_cur, self.dt_cur)
                torch.cuda.synchronize()
                dist.barrier()
                n0 = ops.launch_count
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
                    self._step_body(self.mod_cur, self.dt_cur)
                self.launches_per_step = ops.launch_count - n0
                s._load(keep, None)
            torch.cuda.synchronize()
            dist.barrier()                  # nobody replays / steps before every rank is set up
            self.check_exchange()

    def _step_body(self, mod, dt_dev) -> None:
        s = self.sess
        prec = s.precise
        kw = dict(img_shapes=s.shapes, txt_seq_lens=[s.s_txt], return_dict=False, _mod_override=mod, _out_fp32=prec)
        if self.layout.cfg_degree == 1:
            c, u = self.pipe._predict(s.cur, s.cond, s.uncond, kw, s.enc2, s.x2)
        else:
            pred, _ = self.pipe._predict(s.cur, s.cond, None, kw, None, s.x2)     # x2: [cur; reference-image shard]
            dist.all_gather_into_tensor(self.pair_buf.view(-1), pred.reshape(-1), group=self.pair_group)
            c, u = self.pair_buf[0], self.pair_buf[1]
        if prec:      # fp32 predictions, fp32 master latents updated in place, bf16 copy for the next forward
            torch.ops.qi.cfg_euler_step_f32(c.contiguous(), None if u is None else u.contiguous(), s.cur32, s.cur32,
                                            s.cur, self.scale, dt_dev)
        else:
            torch.ops.qi.cfg_euler_step(c, u, s.cur, s.nxt, self.scale, dt_dev)
            s.cur.copy_(s.nxt)

    def load_inputs(self, cond_h, uncond_h, lat_h, ref_h=None):
        lay = self.layout
        mine = cond_h if (lay.cfg_degree == 1 or lay.branch == 0) else uncond_h
        self.sess.cond.copy_(mine[:, self.t0:self.t1], non_blocking=True)
        if self.sess.uncond is not None:
            self.sess.uncond.copy_(uncond_h[:, self.t0:self.t1], non_blocking=True)
        ref = ref_h[:, self.j0:self.j1] if ref_h is not None else None
        self.sess._load(lat_h[:, self.i0:self.i1], ref)

    def read_latents(self, out_h):
        """This rank's shard of the current latents into its rows of a host buffer (no gather; see `gather_latents`)."""
        out_h[:, self.i0:self.i1].copy_(self.sess.latents, non_blocking=True)

    def gather_latents(self) -> torch.Tensor:
        """Complete current latents [B, S_img, 64] on every rank (device).  The shards of a sequence-parallel group
        are ragged, so each rank drops its shard into a zero buffer and the group sums it (exact: x + 0)."""
        lat = self.sess.latents_master       # fp32 loop state in precise mode
        if self.layout.sp_degree == 1:
            return lat.clone()
        full = torch.zeros(lat.shape[0], self.s_img, lat.shape[2], device=lat.device, dtype=lat.dtype)
        full[:, self.i0:self.i1] = lat
        dist.all_reduce(full, group=self.sp_group)
        return full

    def check_exchange(self) -> None:
        """Raise if a peer flag barrier timed out (a rank died or the launch sequences diverged): after a timeout
        the barriers stop waiting, so every later step would silently read stale peer data."""
        if self.peer is not None and self.peer.timed_out():
            raise RuntimeError("peer-memory exchange: a flag barrier timed out; the latents of this run are invalid")

    # -------------------------------------------------------------- the pipeline call, sharded
    def denoise(self, cond, uncond, latents, height, width, num_inference_steps, true_cfg_scale=4.0,
                ref_tokens=None, ref_hw=None, sigmas=None) -> torch.Tensor:
        """What `B200QwenImagePipeline.__call__` runs when this denoiser is attached: every rank passes the same
        full-size inputs (host or device), the ranks shard them, run the schedule and return the complete final
        latents on every rank.  Setup (peer buffers, modulation table, tuning, graph) is kept across calls of the
        same shapes."""
        if sigmas is not None:
            num_inference_steps = len(sigmas)
        if uncond is None and self.layout.cfg_degree == 2:
            raise ValueError("CFG-parallel layout needs negative_prompt_embeds (or build with cfg_parallel=False)")
        key = (tuple(cond.shape), None if uncond is None else tuple(uncond.shape), tuple(latents.shape), height, width,
               num_inference_steps, float(true_cfg_scale), None if ref_tokens is None else (tuple(ref_tokens.shape), ref_hw),
               None if sigmas is None else tuple(float(v) for v in sigmas), self.tr.state_signature())
        if key != self._setup_key:
            self.setup(cond, uncond, latents, height, width, num_inference_steps, true_cfg_scale,
                       ref_h=ref_tokens, ref_hw=ref_hw, sigmas=sigmas)
            self._setup_key = key
        else:
            self.load_inputs(cond, uncond, latents, ref_tokens)
        for i in range(num_inference_steps):
            self.step(i)
        out = self.gather_latents()
        self.check_exchange()
        return out

    # -------------------------------------------------------------- one step
    def step(self, i: int) -> None:
        from . import ops
        if self.world == 1:
            self.sess.step(i)
            return
        s = self.sess
        if self.graph is not None:
            self.mod_cur.copy_(s.mod_table[i:i + 1])
            self.dt_cur.copy_(s.dts[i:i + 1])
            self.graph.replay()
        else:
            self._step_body(s.mod_table[i:i + 1], s.dts[i:i + 1])

    def exchange_timed_out(self) -> bool:
        """True if a peer barrier gave up waiting (a rank died or the launch sequences diverged)."""
        return self.peer is not None and self.peer.timed_out()

    def close(self):
        self.graph = None
        if self.peer is not None:
            if self.tr._sp is not None:
                self.tr._sp.peer = None
            self.peer.close()
            self.peer = None
        if self.sess is not None:
            self.sess.close()
            self.sess = None
