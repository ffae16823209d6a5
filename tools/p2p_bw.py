"""Peer-write bandwidth over NVLink as the fused exchange kernels see it: SM stores (not the copy engines)
into a CUDA-IPC-mapped region of the other GPU, contiguous and as row segments at the scatter's row stride.
Gives the ceiling the Ulysses scatter epilogues run against.   torchrun --nproc-per-node 2 tools/p2p_bw.py"""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_omni_draft_b200 import _lib  # noqa: E402
from vllm_omni_draft_b200._lib import check  # noqa: E402
from vllm_omni_draft_b200.launcher import init_distributed  # noqa: E402
from vllm_omni_draft_b200.peer import _RawDeviceMemory  # noqa: E402

world, rank, local_rank = init_distributed(int(os.environ.get("WORLD_SIZE", "1")))
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
NBYTES = 256 << 20
lib = _lib.lib()
own = C.c_void_p()
check(lib.qi_peer_alloc(C.byref(own), NBYTES), "qi_peer_alloc")
handle = C.create_string_buffer(64)
check(lib.qi_peer_export(own.value, handle), "qi_peer_export")
handles = [b""] * world
dist.all_gather_object(handles, handle.raw)
peer_rank = (rank + 1) % world
q = C.c_void_p()
check(lib.qi_peer_import(C.create_string_buffer(handles[peer_rank], 64), C.byref(q)), "qi_peer_import")
peer = torch.as_tensor(_RawDeviceMemory(q.value, NBYTES // 2), device=dev).view(torch.bfloat16)
mine = torch.as_tensor(_RawDeviceMemory(own.value, NBYTES // 2), device=dev).view(torch.bfloat16)
src = torch.randn(NBYTES // 2, device=dev).bfloat16()
torch.cuda.synchronize()
dist.barrier()


def timed(fn, nbytes, iters=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    dist.barrier()
    return nbytes / ms * 1e-6      # GB/s


def case(name, dst_of, src_view, nbytes, senders):
    """senders: ranks that write; the others idle (one-directional) or write as well (both directions)."""
    def fn():
        if rank in senders:
            torch.add(src_view, 1, out=dst_of)
    gbs = timed(fn, nbytes)
    vals = [None] * world
    dist.all_gather_object(vals, gbs if rank in senders else None)
    if rank == 0:
        print(f"P2P_BW {name:58s} " + "  ".join(f"r{r}: {v:7.1f} GB/s" for r, v in enumerate(vals) if v is not None), flush=True)


n = 64 << 20                      # 128 MB of bf16 per pass
everyone = list(range(world))
case("contiguous 128 MB, SM stores, one direction", peer[:n], src[:n], 2 * n, [0])
case("contiguous 128 MB, SM stores, all ranks at once", peer[:n], src[:n], 2 * n, everyone)
case("local HBM (same kernel, own region), all ranks", mine[:n], src[:n], 2 * n, everyone)
ld = 3 * 6 * 128                  # row stride of the head owner's [.., 3, H/P, Dh] buffer at P = 4 (elements)
rows = (NBYTES // 2) // ld
for seg in (32, 64, 128):         # 64-, 128-, 256-byte row segments
    dst = peer[:rows * ld].view(rows, ld)[:, :seg]
    s2 = src[:rows * seg].view(rows, seg)
    case(f"{seg * 2:3d}-byte row segments at {ld * 2}-byte stride, all ranks", dst, s2, rows * seg * 2, everyone)
    dstl = mine[:rows * ld].view(rows, ld)[:, :seg]
    case(f"{seg * 2:3d}-byte row segments, local HBM, all ranks", dstl, s2, rows * seg * 2, everyone)
# packet size alone: the same segments at a 256- or 512-byte stride (rows of one head adjacent, as in a
# [.., head, token, Dh] destination layout)
for seg, ld2 in ((32, 128), (64, 128), (64, 256), (128, 256), (128, 128)):
    rows2 = min((NBYTES // 2) // ld2, (64 << 20) // seg)
    dst = peer[:rows2 * ld2].view(rows2, ld2)[:, :seg]
    s2 = src[:rows2 * seg].view(rows2, seg)
    case(f"{seg * 2:3d}-byte row segments at {ld2 * 2}-byte stride, all ranks", dst, s2, rows2 * seg * 2, everyone)
torch.cuda.synchronize()
dist.barrier()
lib.qi_peer_unmap(q.value)
dist.barrier()
lib.qi_peer_free(own.value)
dist.destroy_process_group()
