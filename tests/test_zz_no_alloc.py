"""``i64b_dispatch_plan_static``: the dispatch plan for hosts that cannot allocate device memory.

Who needs it: a serving process whose KV-cache pool already owns all the HBM that was free after loading (a cudaMalloc
made while serving fails there), and an FFI host that is only handed device pointers and has no allocator binding.  For
them the plan's scratch is a fixed 1 MiB array inside ``libi64b.so`` instead of a caller workspace.

What is checked, through the C ABI (ctypes, raw pointers):
  * the five outputs are bit-identical to ``i64b_dispatch_plan``'s (integer work: ``array_equal``, no tolerance) on two
    calls in a row — the second one smaller and differently routed, so anything stale left in the shared scratch by the
    first would show;
  * the calls do not change the device's free memory (nothing was allocated behind the caller's back);
  * a plan beyond the built-in scratch is refused with I64B_ERR_WORKSPACE and a message naming both sizes.

This file is named to be collected last: ``pytest -x`` stops at the first failure, and a failure here must not hide the
parity tests.  It passes alone in a fresh process (the entry point's first call on the device does its own set-up) and
after the rest of the suite (the library is then already warm and the scratch holds nothing useful).
"""
import numpy as np
import pytest
import torch

from vllm_i64 import _C

STATIC_BYTES = 1 << 20          # I64B_DISPATCH_PLAN_STATIC_BYTES, include/i64b.h
ERR_WORKSPACE = -3

needs_gpu = pytest.mark.skipif(not torch.cuda.is_available(), reason="CUDA device required")


def _routes(n, e, k, seed):
    rng = np.random.RandomState(seed)
    r = rng.randint(0, e, size=(k, n)).astype(np.int32)
    r[0, : n // 3] = 1 % e      # a long run to one expert, as in test_dispatch_plan_bit_exact
    return r


def _outputs(k, n, e, dev):
    # filled with a sentinel so an output the kernel skipped cannot pass by accident
    mk = lambda *s: torch.full(s, -7, dtype=torch.int32, device=dev)
    return [mk(k, n), mk(k, e), mk(k, e), mk(k, n), mk(e)]     # scatter, counts, offsets, dest_row, group_rows


def _plan_with_workspace(lib, ids, e, dev):
    k, n = ids.shape
    outs = _outputs(k, n, e, dev)
    ws = torch.empty(lib.i64b_plan_workspace_bytes(n, k, e), dtype=torch.uint8, device=dev)
    _C.check(lib.i64b_dispatch_plan(ids.data_ptr(), k, n, e, *[o.data_ptr() for o in outs], ws.data_ptr(), ws.numel(),
                                    _C.stream_ptr(dev)))
    return outs


def _plan_static(lib, ids, e, outs, dev):
    k, n = ids.shape
    _C.check(lib.i64b_dispatch_plan_static(ids.data_ptr(), k, n, e, *[o.data_ptr() for o in outs], _C.stream_ptr(dev)))


@pytest.mark.gpu
@needs_gpu
def test_static_plan_equals_workspace_plan_twice_in_a_row():
    dev = torch.device("cuda", 0)
    lib = _C.lib()
    # first a multi-tile plan that fills most of the scratch, then a small one with other routes over the same bytes
    cases = [(60001, 8, 2, 11), (1500, 16, 4, 12)]
    ids = [torch.from_numpy(_routes(n, e, k, s)).to(dev) for n, e, k, s in cases]
    want = [_plan_with_workspace(lib, i, c[1], dev) for i, c in zip(ids, cases)]
    got = [_outputs(c[2], c[0], c[1], dev) for c in cases]
    assert lib.i64b_plan_workspace_bytes(cases[0][0], cases[0][2], cases[0][1]) <= STATIC_BYTES
    torch.cuda.synchronize(dev)

    _plan_static(lib, ids[0], cases[0][1], got[0], dev)     # in a fresh process this call also resolves the array's address
    torch.cuda.synchronize(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]
    _plan_static(lib, ids[1], cases[1][1], got[1], dev)     # back to back, same stream, stale scratch underneath
    _plan_static(lib, ids[0], cases[0][1], got[0], dev)     # and the big one again over the small one's leftovers
    torch.cuda.synchronize(dev)
    assert torch.cuda.mem_get_info(dev)[0] == free_before, "the no-allocation entry point changed the device's free memory"

    names = ("scatter_indices", "counts", "offsets", "dest_row", "group_rows")
    for c, w, g in zip(cases, want, got):
        for name, a, b in zip(names, w, g):
            np.testing.assert_array_equal(b.cpu().numpy(), a.cpu().numpy(), err_msg=f"{name} differs for n,e,k,seed={c}")


@pytest.mark.gpu
@needs_gpu
def test_static_plan_nullable_outputs_and_empty_batch():
    dev = torch.device("cuda", 0)
    lib = _C.lib()
    ids = torch.from_numpy(_routes(777, 4, 2, 5)).to(dev)
    want = _plan_with_workspace(lib, ids, 4, dev)
    dest = torch.full((2, 777), -7, dtype=torch.int32, device=dev)
    _C.check(lib.i64b_dispatch_plan_static(ids.data_ptr(), 2, 777, 4, None, None, None, dest.data_ptr(), None,
                                           _C.stream_ptr(dev)))
    np.testing.assert_array_equal(dest.cpu().numpy(), want[3].cpu().numpy())
    counts = torch.full((1, 4), -7, dtype=torch.int32, device=dev)
    _C.check(lib.i64b_dispatch_plan_static(None, 1, 0, 4, None, counts.data_ptr(), None, None, None, _C.stream_ptr(dev)))
    assert counts.cpu().tolist() == [[0, 0, 0, 0]]


def test_static_plan_refuses_what_does_not_fit():
    """Needs no GPU: the size check comes before the first CUDA call."""
    lib = _C.lib()
    n, k, e = 200000, 2, 8
    need = lib.i64b_plan_workspace_bytes(n, k, e)
    assert need > STATIC_BYTES
    assert lib.i64b_dispatch_plan_static(None, k, n, e, None, None, None, None, None, None) == -1      # null ids first
    fake = 256                                                # never dereferenced: the call is refused on size alone
    assert lib.i64b_dispatch_plan_static(fake, k, n, e, None, None, None, None, None, None) == ERR_WORKSPACE
    msg = lib.i64b_last_error().decode()
    assert str(need) in msg and str(STATIC_BYTES) in msg and "i64b_dispatch_plan" in msg
    assert lib.i64b_dispatch_plan_static(fake, 5, 10, e, None, None, None, None, None, None) == -1     # top_k > 4
