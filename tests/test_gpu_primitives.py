"""GPU parity of the small primitives (cornerstone_octree_b200/csrc/primitives.cu) through the C ABI against the numpy
restatement in oracle/primitives_np.py (integer / byte work: bit-exact; float work: same element-wise expression)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _t(a):
    import torch
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    elif a.dtype == np.uint64:
        a = a.view(np.int64)
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _n(t, dtype=None):
    a = t.cpu().numpy()
    return a.view(dtype) if dtype is not None else a


@pytest.mark.parametrize("n", [0, 1, 1000, 1 << 20])
def test_fill_scale_increment(n):
    import torch
    from cornerstone_octree_b200 import ops
    from oracle import primitives_np as P
    rng = np.random.default_rng(n + 1)
    for dtype, val in ((np.int32, -7), (np.int64, 1 << 40), (np.float32, 1.5), (np.float64, -2.25), (np.uint8, 3)):
        t = torch.zeros(n, dtype=getattr(torch, np.dtype(dtype).name), device="cuda")
        assert np.array_equal(_n(ops.fill(t, val)), P.fill(n, val, dtype))
    for dtype in (np.float32, np.float64):
        a = rng.standard_normal(n).astype(dtype)
        assert np.array_equal(_n(ops.scale(_t(a), 2.0 * 1.3)), P.scale(a, 2.0 * 1.3))
    for dtype, val in ((np.uint32, 0xfffffff0), (np.uint64, (1 << 64) - 5)):
        a = rng.integers(0, np.iinfo(dtype).max, n, dtype=dtype)
        assert np.array_equal(_n(ops.increment(_t(a), val), dtype), P.increment(a, val))


@pytest.mark.parametrize("n", [0, 1, 31, 100003, 3_000_000])
def test_reductions(n):
    from cornerstone_octree_b200 import ops
    from oracle import primitives_np as P
    rng = np.random.default_rng(n + 7)
    for dtype in (np.float32, np.float64):
        x, y, z = (rng.standard_normal(n).astype(dtype) for _ in range(3))
        assert ops.max_norm_square(_t(x), _t(y), _t(z)) == P.max_norm_square(x, y, z)
    a = rng.integers(0, 1 << 32, n, dtype=np.uint32)
    assert ops.reduce_sum(_t(a), init=5) == P.reduce_sum(a, 5)
    b = rng.integers(0, 4, n, dtype=np.uint32)
    assert ops.count(_t(b), 1) == P.count(b, 1)
    c = rng.integers(0, 3, n, dtype=np.uint64) + np.uint64(1 << 63)
    assert ops.count(_t(c), (1 << 63) + 2) == P.count(c, (1 << 63) + 2)


@pytest.mark.parametrize("kdtype", [np.uint32, np.uint64])
def test_lower_bound(kdtype):
    from cornerstone_octree_b200 import ops
    from oracle import primitives_np as P
    rng = np.random.default_rng(3)
    hi = np.iinfo(kdtype).max
    a = np.sort(rng.integers(0, hi, 200001, dtype=kdtype))
    a[1000:1010] = a[1000]  # duplicates
    vals = np.concatenate([rng.integers(0, hi, 5000, dtype=kdtype), a[::97], np.array([0, hi], dtype=kdtype)])
    for ib, idt in ((32, np.uint32), (64, np.uint64)):
        got = _n(ops.lower_bound_many(_t(a), _t(vals), ib), idt)
        assert np.array_equal(got, P.lower_bound(a, vals).astype(idt))
    for v in (0, int(a[1000]), int(a[-1]), hi):
        assert ops.lower_bound(_t(a), v) == int(P.lower_bound(a, kdtype(v)))
    assert ops.lower_bound(_t(a[:0]), 5) == 0
    s = np.sort(rng.integers(-1000, 1000, 999).astype(np.int32))
    assert ops.lower_bound(_t(s), -3, unsigned=False) == int(P.lower_bound(s, np.int32(-3)))
    f = np.sort(rng.standard_normal(999).astype(np.float32))
    assert ops.lower_bound(_t(f), 0.25) == int(P.lower_bound(f, np.float32(0.25)))


@pytest.mark.parametrize("mean_len", [0.5, 15, 400])
def test_segment_max(mean_len):
    import torch
    from cornerstone_octree_b200 import ops
    from oracle import primitives_np as P
    rng = np.random.default_rng(int(mean_len * 10))
    nseg = 3000
    lens = rng.poisson(mean_len, nseg)
    for sdt in (np.uint32, np.uint64):
        seg = np.concatenate([[0], np.cumsum(lens)]).astype(sdt)
        for idt, odt in ((np.float32, np.float32), (np.float64, np.float32), (np.float64, np.float64)):
            h = (rng.random(int(seg[-1])) - 0.1).astype(idt)  # some negative values: the maximum starts at 0
            got = _n(ops.segment_max(_t(h), _t(seg), getattr(torch, np.dtype(odt).name)))
            assert np.array_equal(got, P.segment_max(h, seg, odt))


@pytest.mark.parametrize("kdtype", [np.uint32, np.uint64])
def test_sort_keys(kdtype):
    from cornerstone_octree_b200 import ops
    from oracle import primitives_np as P
    rng = np.random.default_rng(11)
    for n in (0, 1, 4097, 1_000_003):
        a = rng.integers(0, np.iinfo(kdtype).max, n, dtype=kdtype)
        assert np.array_equal(_n(ops.sort_keys(_t(a)), kdtype), P.sort_keys(a))


@pytest.mark.parametrize("kdtype", [np.uint32, np.uint64])
def test_sfc_gaps(oracle, kdtype):
    """countSfcGaps + exclusive scan + fillSfcGaps == computeSpanningTree; the reference's golden vectors for
    spanSfcRange (test/unit/sfc/common.cpp:376-419) as single intervals"""
    import torch
    from cornerstone_octree_b200 import ops
    from oracle import primitives_np as P
    ml = 10 if kdtype == np.uint32 else 21
    top = 1 << (3 * ml)
    rng = np.random.default_rng(5)
    inner = np.unique(rng.integers(1, top, 500, dtype=np.uint64) & ~np.uint64(0o7))
    cases = [np.array([0, top], dtype=kdtype), np.array([0, 0o7777, top], dtype=kdtype),
             np.array([0, 1, 0o742, 0o41305, 0o46000, top], dtype=kdtype),
             np.concatenate([[0], inner, [top]]).astype(kdtype)]
    for tree in cases:
        ops_ref, new_ref = P.count_fill_sfc_gaps(oracle, tree)
        ops_gpu = ops.count_sfc_gaps(_t(tree))
        assert np.array_equal(_n(ops_gpu)[:-1], ops_ref)
        scanned = torch.zeros_like(ops_gpu)
        scanned[1:] = torch.cumsum(ops_gpu[:-1], 0)
        new_gpu = ops.fill_sfc_gaps(_t(tree), scanned, int(ops_ref.sum()))
        assert np.array_equal(_n(new_gpu, kdtype), new_ref)


def test_gather_ranges():
    from cornerstone_octree_b200 import ops
    from oracle import primitives_np as P
    rng = np.random.default_rng(9)
    n = 500000
    for idt in (np.uint32, np.uint64):
        starts = np.sort(rng.choice(n - 200, 700, replace=False)).astype(idt)
        lens = rng.integers(1, 150, starts.size).astype(idt)
        scan = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(idt)
        total = int(lens.sum())
        for dtype, cols in ((np.float32, 1), (np.float64, 1), (np.float32, 3), (np.float32, 4), (np.int32, 1)):
            src = rng.integers(-1000, 1000, (n, cols)).astype(dtype)
            if cols == 1:
                src = src.reshape(n)
            got = _n(ops.gather_ranges(_t(scan), _t(starts), _t(src), total))
            assert np.array_equal(got, P.gather_ranges(scan, starts, src, total))
