"""cstone_b200_exclusive_scan_u32_noalloc (no counterpart in the reference):
the scan whose scratch is a file-scope device array inside the library, for callers that cannot call
cudaMalloc.  Called through the C ABI and compared bit-for-bit (integer work) with the existing entry point
cstone_b200_exclusive_scan_u32 on two calls in a row; the shared array keeps the first call's contents, so the second
call (fewer tiles, different data) proves that what must start at zero is cleared.  Named test_zz_* so that pytest
collects it last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_N = 16777216  # CSTONE_B200_SCAN_NOALLOC_MAX_N, include/cstone_b200.h


def _scan_both(values, init):
    """(existing entry point with caller scratch, no-alloc entry point) on the same device input"""
    import torch
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200._lib import call
    n = values.numel()
    # every buffer exists before the call under test: it must not allocate, so it gets none of torch's help either
    ref_out = torch.empty_like(values)
    new_out = torch.empty_like(values)
    temp = torch.empty(int(call("scan_temp_bytes", n)) + 256, dtype=torch.uint8, device="cuda")
    call("exclusive_scan_u32", ops._p(values), ops._p(ref_out), n, init, ops._p(temp), ops._stream())
    call("exclusive_scan_u32_noalloc", ops._p(values), ops._p(new_out), n, init, ops._stream())
    torch.cuda.synchronize()
    return ref_out, new_out


def test_noalloc_scan_matches_existing_entry_point_twice():
    import torch
    rng = np.random.default_rng(2024)
    # first call: many tiles (fills ticket + status words); second: fewer tiles and other data on the SAME scratch;
    # third: the stated maximum
    for n, init in ((3_000_001, 0), (70_001, 12345), (MAX_N, 7)):
        host = rng.integers(0, 200, size=n, dtype=np.int64).astype(np.uint32)
        values = torch.from_numpy(host.view(np.int32)).cuda()
        ref_out, new_out = _scan_both(values, init)
        assert torch.equal(ref_out, new_out), "n = %d" % n
        # and both agree with the definition on the host
        expect = (np.cumsum(host, dtype=np.uint64) - host + init).astype(np.uint32)
        assert np.array_equal(new_out.cpu().numpy().view(np.uint32), expect)


def test_noalloc_scan_in_place_and_empty():
    import torch
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200._lib import call
    host = np.arange(1, 5001, dtype=np.uint32)
    values = torch.from_numpy(host.view(np.int32)).cuda()
    call("exclusive_scan_u32_noalloc", ops._p(values), ops._p(values), values.numel(), 0, ops._stream())  # out aliases in
    torch.cuda.synchronize()
    assert np.array_equal(values.cpu().numpy().view(np.uint32), (np.cumsum(host, dtype=np.uint64) - host).astype(np.uint32))
    call("exclusive_scan_u32_noalloc", None, None, 0, 0, ops._stream())  # n = 0: nothing to do, no error


def test_noalloc_scan_rejects_sizes_beyond_the_maximum():
    import torch
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200._lib import CstoneError, call, launch_count
    buf = torch.zeros(16, dtype=torch.int32, device="cuda")  # never touched: the size check comes first
    before = launch_count()
    with pytest.raises(CstoneError) as err:
        call("exclusive_scan_u32_noalloc", ops._p(buf), ops._p(buf), MAX_N + 1, 0, ops._stream())
    assert err.value.status == -3  # CSTONE_E_TOO_LARGE
    assert launch_count() == before
