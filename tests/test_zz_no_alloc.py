"""b200ir_l1_loss_fwd_noalloc — the allocation-free variant of b200ir_l1_loss_fwd (scratch in a file-scope __device__ array
of the library instead of a caller buffer; include/b200ir.h).  It has no counterpart in the reference,
so the check is against the existing entry point: bitwise, since both run the same
fixed-order partial sums when given the same number of parts, and against torch's mean |a - b| within the tolerance the
other L1 tests use.  Called through the C-ABI (ctypes), twice in a row with different inputs, because every call shares
the one scratch array and it keeps the previous call's contents.

This file is named to be collected last.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

ELEMS_PER_PART = 4096            # B200IR_L1_NOALLOC_ELEMS_PER_PART
MAX_N = 4096 * 4096              # B200IR_L1_NOALLOC_MAX_N


def _pair(shape, seed, dtype):
    g = torch.Generator().manual_seed(seed)
    a = torch.rand(shape, generator=g).cuda().to(dtype).contiguous()
    b = torch.rand(shape, generator=g).cuda().contiguous()
    return a, b


def _both(a, b):
    from b200ir import _lib as L
    from b200ir.ops import _DT, _stream
    n = a.numel()
    nparts = (n + ELEMS_PER_PART - 1) // ELEMS_PER_PART
    parts = torch.empty((nparts,), device="cuda", dtype=torch.float32)
    want = torch.full((), -1.0, device="cuda", dtype=torch.float32)
    got = torch.full((), -2.0, device="cuda", dtype=torch.float32)
    L.check(L.lib.b200ir_l1_loss_fwd(a.data_ptr(), _DT[a.dtype], b.data_ptr(), _DT[b.dtype], n, parts.data_ptr(), nparts,
                                     want.data_ptr(), _stream()), "b200ir_l1_loss_fwd")
    L.check(L.lib.b200ir_l1_loss_fwd_noalloc(a.data_ptr(), _DT[a.dtype], b.data_ptr(), _DT[b.dtype], n, got.data_ptr(),
                                             _stream()), "b200ir_l1_loss_fwd_noalloc")
    torch.cuda.synchronize()
    return got.item(), want.item()


def test_l1_loss_noalloc_matches_existing_entry_point_twice():
    # call 1: a large batch (576 parts), call 2: a small ragged one (2 parts) whose partial sums overwrite only the head
    # of the shared array, call 3: the first shape again with other data and a bf16 prediction
    cases = [((8, 3, 128, 128 + 64), 1, torch.float32), ((1, 3, 37, 53), 2, torch.float32),
             ((8, 3, 128, 128 + 64), 3, torch.bfloat16)]
    for shape, seed, dtype in cases:
        a, b = _pair(shape, seed, dtype)
        got, want = _both(a, b)
        assert got == want, (shape, got, want)                       # same fixed-order sums: bitwise
        ref = (a.double() - b.double()).abs().mean().item()
        assert abs(got - ref) <= 1e-5 * abs(ref), (shape, got, ref)


def test_l1_loss_noalloc_allocates_nothing_and_rejects_oversize():
    from b200ir import _lib as L
    from b200ir.ops import _DT, _stream
    a, b = _pair((2, 3, 64, 64), 7, torch.float32)
    out = torch.empty((), device="cuda", dtype=torch.float32)
    _both(a, b)                                                       # set-up (symbol address look-up) done
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    stats0 = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    for _ in range(3):
        L.check(L.lib.b200ir_l1_loss_fwd_noalloc(a.data_ptr(), _DT[a.dtype], b.data_ptr(), _DT[b.dtype], a.numel(),
                                                 out.data_ptr(), _stream()), "b200ir_l1_loss_fwd_noalloc")
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == stats0
    assert torch.cuda.mem_get_info()[0] == free0
    # beyond the built-in scratch: refused before any launch (the pointers are never dereferenced), clear message
    rc = L.lib.b200ir_l1_loss_fwd_noalloc(a.data_ptr(), _DT[a.dtype], b.data_ptr(), _DT[b.dtype], MAX_N + 1,
                                          out.data_ptr(), _stream())
    assert rc == -3, rc                                               # B200IR_E_WORKSPACE
    msg = L.lib.b200ir_last_error_string().decode()
    assert "exceeds the built-in scratch" in msg and str(MAX_N) in msg, msg
