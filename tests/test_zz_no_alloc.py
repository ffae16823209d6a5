"""qi_peer_barrier_static: the flag barrier for hosts that cannot allocate device memory.

Collected last on purpose (the file name sorts after every other test file): the driver runs pytest with -x, and a
failure here should not hide the parity tests.

Who needs this entry point: a vLLM worker that has already given the rest of HBM to the engine's block pool, a step
loop that cannot take the device-wide stall of cudaMalloc / cudaFree, or a non-Python host bound to the C ABI that
has no device allocator at all.  qi_peer_barrier asks such a caller for two scratch buffers (the device table of
signal-array pointers and the {epoch, timeout flag} words); qi_peer_barrier_static keeps both in a fixed 256-byte
array inside libqi_b200.so.  The signal arrays themselves remain the caller's, as with the allocating entry point.

The tests call the C ABI directly through ctypes and compare with qi_peer_barrier on the same sequence of calls.
The shared array survives from call to call (and from test to test in one process), so every sequence starts with
restart=1, and the checks hold both in a fresh process and after the rest of the suite has run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from vllm_omni_draft_b200 import _lib as L
    return L


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _static_barrier(lib, sig, n, me, restart, state_out, timeout_ms=2000):
    """One call of the new entry point; `sig` is the list of signal arrays (device tensors), passed as a HOST table."""
    table = (C.c_void_p * len(sig))(*[t.data_ptr() for t in sig])   # n is passed as given (the bad-argument cases)
    return lib.qi_peer_barrier_static(table, n, me, restart, state_out.data_ptr() if state_out is not None else None,
                                      timeout_ms, _stream())


def test_static_barrier_matches_allocating_barrier_two_calls_in_a_row(cuda_device):
    """Two barriers back to back through each entry point, single-rank group (a rank signals itself): after every
    call the signal array and the {epoch, timeout flag} words must be equal, bit for bit (integer work, so the
    tolerance is zero, as in test_peer_barrier_single_rank).  Then a new count on a fresh signal array: restart=1
    must bring the shared epoch back to zero, otherwise the barrier would wait for a value the array never reaches
    and the timeout flag would come back set."""
    L = _lib()
    lib = L.lib()
    # what the test allocates here are the caller's inputs/outputs in both cases; the scratch differs
    sig_ref = torch.zeros(8, dtype=torch.int32, device="cuda")
    table_ref = torch.tensor([sig_ref.data_ptr()], dtype=torch.int64, device="cuda")   # scratch of the old entry
    state_ref = torch.zeros(2, dtype=torch.int32, device="cuda")                        # scratch of the old entry
    sig_new = torch.zeros(8, dtype=torch.int32, device="cuda")
    state_new = torch.full((2,), -1, dtype=torch.int32, device="cuda")                  # output only
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for call in range(2):
        L.check(lib.qi_peer_barrier(table_ref.data_ptr(), 1, 0, state_ref.data_ptr(), 2000, _stream()),
                "qi_peer_barrier")
        L.check(_static_barrier(lib, [sig_new], 1, 0, restart=int(call == 0), state_out=state_new),
                "qi_peer_barrier_static")
        torch.cuda.synchronize()
        assert torch.equal(sig_new, sig_ref), (call, sig_new.tolist(), sig_ref.tolist())
        assert torch.equal(state_new, state_ref), (call, state_new.tolist(), state_ref.tolist())
    stop.record()
    torch.cuda.synchronize()
    assert sig_new.tolist()[0] == 2 and state_new.tolist() == [2, 0]
    assert start.elapsed_time(stop) < 1000.0, "a barrier spun until its 2 s timeout"

    # a second group's worth of fresh arrays: the shared epoch (2 by now) must not leak into the new count
    sig_ref.zero_(), state_ref.zero_()
    sig_new2 = torch.zeros(8, dtype=torch.int32, device="cuda")
    for call in range(2):
        L.check(lib.qi_peer_barrier(table_ref.data_ptr(), 1, 0, state_ref.data_ptr(), 2000, _stream()),
                "qi_peer_barrier")
        L.check(_static_barrier(lib, [sig_new2], 1, 0, restart=int(call == 0), state_out=state_new),
                "qi_peer_barrier_static")
    torch.cuda.synchronize()
    assert torch.equal(sig_new2, sig_ref) and torch.equal(state_new, state_ref)
    assert state_new.tolist() == [2, 0]
    assert sig_new.tolist()[0] == 2, "the first group's array was written after its table entry was replaced"


def test_static_barrier_without_state_output_and_epoch_carry_over(cuda_device):
    """state_out_dev may be NULL; the epoch carries over between calls that do not restart (third call sees 3)."""
    L = _lib()
    lib = L.lib()
    sig = torch.zeros(8, dtype=torch.int32, device="cuda")
    state = torch.full((2,), -1, dtype=torch.int32, device="cuda")
    L.check(_static_barrier(lib, [sig], 1, 0, restart=1, state_out=None), "qi_peer_barrier_static")
    L.check(_static_barrier(lib, [sig], 1, 0, restart=0, state_out=None), "qi_peer_barrier_static")
    L.check(_static_barrier(lib, [sig], 1, 0, restart=0, state_out=state), "qi_peer_barrier_static")
    torch.cuda.synchronize()
    assert int(sig[0]) == 3 and state.tolist() == [3, 0]


def test_static_barrier_rejects_what_does_not_fit(cuda_device):
    """More peers than the fixed table holds is an argument error with a message, not a silent overflow; nothing is
    launched for it (the signal array stays untouched)."""
    L = _lib()
    lib = L.lib()
    sig = torch.zeros(16, dtype=torch.int32, device="cuda")
    rc = _static_barrier(lib, [sig] * 9, 9, 0, restart=1, state_out=None)
    assert rc == -1
    msg = lib.qi_last_cuda_error().decode()
    assert "n_peers = 9" in msg and "8 entries" in msg
    with pytest.raises(RuntimeError, match="invalid argument"):
        L.check(rc, "qi_peer_barrier_static")
    assert _static_barrier(lib, [sig], 0, 0, restart=1, state_out=None) == -1
    assert _static_barrier(lib, [sig, sig], 2, 2, restart=1, state_out=None) == -1
    assert lib.qi_peer_barrier_static(None, 1, 0, 1, None, 2000, _stream()) == -1
    torch.cuda.synchronize()
    assert int(sig.abs().sum()) == 0
