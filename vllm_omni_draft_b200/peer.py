"""Peer-memory exchange buffers of one Ulysses sequence-parallel group (one process per GPU).

The reference has no parallelism.  SURVEY §8e asks for the head<->sequence
all-to-all around attention; here that exchange is folded into the kernels on either side of it:

* the QKV GEMM's epilogue (or ``qi_qknorm_rope_scatter``, K2) stores every q/k/v head vector straight into
  the HBM of the rank that owns the head — into that rank's ``recv`` buffer ``[B, 3, H/P, S_total, Dh]``;
* ``qi_attention_fwd_scatter`` (K4) stores every output row straight into the HBM of the rank that
  owns the token — into that rank's ``back`` buffer ``[P·H/P, B, S_loc_max, Dh]`` (one K-chunk of Dh columns
  per (source rank, head)), which the out-projection GEMM reads as a K-chunked A operand;
* both buffers are HEAD-MAJOR: the rows of one head are adjacent at the destination, because the peer
  write rate depends on the destination stride (tools/p2p_bw.py: 128-byte segments 574 GB/s at a 256-byte
  row stride, 324 GB/s at the 4608-byte stride of a token-major ``[.., 3, H/P, Dh]`` row);
* between producer and consumer sits ``qi_peer_barrier`` (a 32-thread flag kernel, overlapped with its
  neighbours' tails by programmatic dependent launch), twice per block.  ``QI_PEER_FOLD=1`` folds the flags into
  the kernels instead (``qi_peer_signal_t``: the producer's last CTA release-increments this rank's flag on every
  peer, the consumer's loader warp acquire-polls its own flags before the first load; ``sig_qkv`` rides in
  ``dst_qkv.notify`` / ``dst_o.wait``, ``sig_o`` in ``dst_o.notify`` and the out-projection GEMM's ``wait``).
  Measured on 4 GPUs (cfg-2, CFG x Ulysses-2, interleaved runs on one box) the folded form is 1.7 % SLOWER
  (39.97 vs 39.27 ms/step): with PDL the barrier launches cost nothing that the in-kernel poll does not also
  cost, and 216 attention CTAs polling system-scope flags is not free.  It stays as a tested option.

Host layer only: allocate one peer-mappable region per rank (``qi_peer_alloc``), ship the 64-byte
CUDA-IPC handles over ``torch.distributed`` and map the peers' regions (``qi_peer_import``).  No
data moves through this module.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Sequence

import torch
import torch.distributed as dist

from . import _lib, ops
from ._lib import PeerScatter, PeerSignal, check

_ALIGN = 1024
_SIG_BYTES = 1024


def _round_up(n: int, a: int = _ALIGN) -> int:
    return (n + a - 1) // a * a


class _RawDeviceMemory:
    """CUDA array interface over a raw device pointer (the memory is owned by PeerExchange)."""

    def __init__(self, ptr: int, n_int16: int):
        self.__cuda_array_interface__ = {"shape": (n_int16,), "typestr": "<i2", "data": (ptr, False),
                                         "version": 3, "strides": None}


def shard_counts(n: int, parts: int) -> List[int]:
    """Ragged split of n rows over `parts` ranks: the first n % parts ranks hold one extra row."""
    q, r = divmod(n, parts)
    return [q + (1 if p < r else 0) for p in range(parts)]


def prefix(counts: Sequence[int]) -> List[int]:
    out = [0]
    for c in counts:
        out.append(out[-1] + int(c))
    return out


class PeerExchange:
    def __init__(self, group, size: int, rank: int, device: torch.device, heads: int, head_dim: int,
                 batch: int, rows_per_rank: Sequence[int], timeout_ms: int = 2000):
        if heads % size != 0:
            raise ValueError("num_attention_heads must be divisible by the sequence-parallel degree")
        if size > 8:
            raise ValueError("at most 8 ranks per sequence-parallel group (one NVSwitch node)")
        self.group, self.size, self.rank, self.device = group, size, rank, device
        self.heads, self.head_dim, self.batch = heads, head_dim, batch
        self.hs = heads // size
        self.counts = [int(c) for c in rows_per_rank]
        self.pre = prefix(self.counts)
        self.s_total, self.s_max, self.s_loc = self.pre[-1], max(self.counts), self.counts[rank]
        self.timeout_ms = timeout_ms
        row = self.hs * head_dim
        recv_bytes = batch * self.s_total * 3 * row * 2
        back_bytes = size * batch * self.s_max * row * 2
        self._recv_off = _SIG_BYTES
        self._back_off = self._recv_off + _round_up(recv_bytes)
        self.nbytes = self._back_off + _round_up(back_bytes)
        lib = _lib.lib()
        with torch.cuda.device(device):
            p = C.c_void_p()
            check(lib.qi_peer_alloc(C.byref(p), self.nbytes), "qi_peer_alloc")
            self._own = p.value
            handle = C.create_string_buffer(64)
            check(lib.qi_peer_export(self._own, handle), "qi_peer_export")
            handles: List[bytes] = [b""] * size
            dist.all_gather_object(handles, handle.raw, group=group)
            self.ptrs: List[int] = []
            self._mapped: List[int] = []
            for r in range(size):
                if r == rank:
                    self.ptrs.append(self._own)
                else:
                    q = C.c_void_p()
                    check(lib.qi_peer_import(C.create_string_buffer(handles[r], 64), C.byref(q)), "qi_peer_import")
                    self.ptrs.append(q.value)
                    self._mapped.append(q.value)
        self.sig_table = torch.tensor(self.ptrs, dtype=torch.int64, device=device)   # signal array at offset 0
        self.state = torch.zeros(2, dtype=torch.int32, device=device)
        # folded flags: slot arrays of the two exchanges at byte offsets 64 and 128 of every rank's region
        self.fold = os.environ.get("QI_PEER_FOLD", "0") == "1"
        self._slot_tables = [torch.tensor([q + off for q in self.ptrs], dtype=torch.int64, device=device)
                             for off in (64, 128)]
        self.sig_state = torch.zeros(2, 4, dtype=torch.int32, device=device)   # per exchange: epoch, CTA count, timeout
        self.sig_qkv, self.sig_o = PeerSignal(), PeerSignal()
        for i, sg in enumerate((self.sig_qkv, self.sig_o)):
            sg.slots, sg.state = self._slot_tables[i].data_ptr(), self.sig_state[i].data_ptr()
            sg.n_peers, sg.my_index, sg.timeout_ms = size, rank, int(timeout_ms)
        # local views (bf16) of this rank's own region
        raw = torch.as_tensor(_RawDeviceMemory(self._own, self.nbytes // 2), device=device).view(torch.bfloat16)
        self._raw = raw
        r0 = self._recv_off // 2
        self.recv = raw[r0:r0 + recv_bytes // 2].view(batch, 3, self.hs, self.s_total, head_dim)
        b0 = self._back_off // 2
        self.back = raw[b0:b0 + back_bytes // 2].view(size * self.hs, batch, self.s_max, head_dim)
        # scatter descriptors
        self.dst_qkv = PeerScatter()
        self.dst_o = PeerScatter()
        for d, off in ((self.dst_qkv, self._recv_off), (self.dst_o, self._back_off)):
            d.n_peers, d.my_index = size, rank
            for r in range(size):
                d.base[r] = self.ptrs[r] + off
            for i, v in enumerate(self.pre):
                d.row_prefix[i] = v
        self.dst_o.chunk_stride = batch * self.s_max * row          # one source rank's H/P head chunks
        # the same descriptors without flags (barrier-kernel protocol)
        self.dst_qkv_plain = PeerScatter.from_buffer_copy(self.dst_qkv)
        self.dst_o_plain = PeerScatter.from_buffer_copy(self.dst_o)
        if self.fold:
            self.dst_qkv.notify = C.addressof(self.sig_qkv)
            self.dst_o.wait = C.addressof(self.sig_qkv)
            self.dst_o.notify = C.addressof(self.sig_o)
        self.o_ld = head_dim
        self.o_batch_stride = self.s_max * head_dim
        self.o_head_stride = batch * self.s_max * head_dim
        torch.cuda.synchronize(device)
        dist.barrier(group=group)      # every region is zeroed and mapped before the first flag is written

    def barrier(self) -> None:
        ops.peer_barrier(self.sig_table, self.size, self.rank, self.state, self.timeout_ms)

    def timed_out(self) -> bool:
        return bool(self.state[1].item() != 0) or bool((self.sig_state[:, 2] != 0).any().item())

    def close(self) -> None:
        if getattr(self, "_own", None) is None:
            return
        lib = _lib.lib()
        torch.cuda.synchronize(self.device)
        try:
            dist.barrier(group=self.group)     # nobody unmaps while a peer may still store
        except Exception:                      # pragma: no cover - best effort at teardown
            pass
        self.recv = self.back = self._raw = None
        for p in self._mapped:
            lib.qi_peer_unmap(p)
        lib.qi_peer_free(self._own)
        self._own = None
        self._mapped = []
