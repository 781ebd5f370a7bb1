"""CPU-side checks of the domain engine's boundary: the library exports every symbol include/cstone_b200_domain.h
declares, and the host part of the decomposition (uniformBins / makeSfcAssignment / limitBoundaryShifts /
limitBoxShrinking) reproduces the reference's golden vectors [ref: test/unit/domain/domaindecomp.cpp:39-153,
test/unit/sfc/box.cpp] and the reference build on random trees.  No GPU, no compute calls on the device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UMAX = 0xFFFFFFFF


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _lib():
    import cornerstone_octree_b200 as pkg
    return pkg._lib.lib


def uniform_bins(counts, nbins, lib=None, name="cstone_b200_host_uniform_bins"):
    lib = lib or _lib()
    counts = np.ascontiguousarray(counts, dtype=np.uint32)
    bins = np.zeros(nbins + 1, dtype=np.int32)
    bc = np.zeros(nbins, dtype=np.uint32)
    getattr(lib, name)(_p(counts), C.c_int32(counts.size), C.c_int32(nbins), _p(bins), _p(bc))
    return bins, bc


def assignment(nranks, leaves, counts, old=None, lib=None, prefix="cstone_b200_host_"):
    lib = lib or _lib()
    ks = "k32" if leaves.dtype == np.uint32 else "k64"
    counts = np.ascontiguousarray(counts, dtype=np.uint32)
    b = np.zeros(nranks + 1, dtype=leaves.dtype)
    rc = np.zeros(nranks, dtype=np.uint32)
    f = getattr(lib, prefix + "assignment_" + ks)
    f(C.c_int(nranks), _p(leaves), _p(counts), C.c_int32(counts.size), _p(old), _p(b), _p(rc))
    return b, rc


def limit_shifts(nranks, old, new, new_counts, leaves, counts, lib=None, prefix="cstone_b200_host_"):
    lib = lib or _lib()
    ks = "k32" if leaves.dtype == np.uint32 else "k64"
    counts = np.ascontiguousarray(counts, dtype=np.uint32)
    b = np.array(new, dtype=leaves.dtype)
    rc = np.array(new_counts, dtype=np.uint32)
    old = np.array(old, dtype=leaves.dtype)
    f = getattr(lib, prefix + "limit_boundary_shifts_" + ks)
    f(C.c_int(nranks), _p(old), _p(b), _p(leaves), _p(counts), C.c_int32(counts.size), _p(rc))
    return b, rc


def test_domain_exports_match_header():
    import cornerstone_octree_b200 as pkg
    header = open(os.path.join(ROOT, "include", "cstone_b200_domain.h")).read()
    declared = set(re.findall(r"\b(cstone_b200_[a-z0-9_]+)\s*\(", header))
    for stem in re.findall(r"\b(cstone_b200_domain_[a-z_]+_)##KS##_##TS", header):
        for ks in ("k32", "k64"):
            for ts in ("f32", "f64"):
                declared.add("%s%s_%s" % (stem, ks, ts))
    assert len(declared) >= 4 * 12 + 10
    lib = C.CDLL(pkg.LIB_PATH)
    assert [n for n in sorted(declared) if not hasattr(lib, n)] == []
    bound = {"cstone_b200_" + n for n in pkg._lib.DOMAIN_EXPORTED}
    assert declared - bound == set()


def test_uniform_bins_golden():
    """[ref: test/unit/domain/domaindecomp.cpp:39-88]"""
    bins, bc = uniform_bins([UMAX - 10, 5, 5, UMAX - 11, 5, 6], 2)
    assert bins.tolist() == [0, 3, 6] and bc.tolist() == [UMAX, UMAX]
    bins, _ = uniform_bins([5, 5, 5, 15, 1, 0], 2)
    assert bins.tolist() == [0, 3, 6]
    _, bc = uniform_bins([15, 0, 1, 5, 5, 5], 2)
    assert bc.min() == 15 and bc.max() == 16
    _, bc = uniform_bins([4, 3, 4, 3, 4, 3, 4, 3, 4, 3], 7)
    assert bc.min() == 3 and bc.max() == 7


def test_make_assignment_golden():
    """[ref: test/unit/domain/domaindecomp.cpp:90-101]"""
    b, rc = assignment(2, np.array([0, 10, 20, 30, 40], dtype=np.uint64), [5, 5, 5, 5])
    assert b.tolist() == [0, 20, 40] and rc.tolist() == [10, 10]


def test_limit_boundary_shifts_golden():
    """[ref: test/unit/domain/domaindecomp.cpp:119-153]"""
    leaves = np.array([0, 5, 10, 15, 20, 25, 30], dtype=np.uint64)
    counts = [1, 2, 3, 4, 5, 6]
    probe = [0, 10, 20, 30]
    b, rc = limit_shifts(3, probe, [0, 25, 30, 30], [15, 6, 0], leaves, counts)
    assert b.tolist() == [0, 20, 30, 30] and rc.tolist() == [10, 11, 0]
    # nothing to limit: boundaries and counts stay as they were
    b, rc = limit_shifts(3, probe, probe, [3, 7, 11], leaves, counts)
    assert b.tolist() == probe and rc.tolist() == [3, 7, 11]


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
def test_assignment_vs_reference(ref, port, kd):
    import gen
    rng = np.random.default_rng(7)
    for trial, nranks in enumerate((1, 2, 3, 8, 13, 64)):
        keys = np.sort(gen.gaussian_keys(20000 + 977 * trial, kd, seed=trial))
        leaves, counts = port.compute_octree(keys, 16 + 5 * trial)
        b, rc = assignment(nranks, leaves, counts)
        rb, rrc = assignment(nranks, leaves, counts, lib=ref.lib, prefix="cstone_ref_")
        assert np.array_equal(b, rb) and np.array_equal(rc, rrc)
        ub, ubc = uniform_bins(counts, nranks)
        rub, rubc = uniform_bins(counts, nranks, lib=ref.lib, name="cstone_ref_uniform_bins")
        assert np.array_equal(ub, rub) and np.array_equal(ubc, rubc)
        assert int(rc.sum()) == keys.size and b[0] == leaves[0] and b[-1] == leaves[-1]
        # a second, shifted particle set with the old boundaries limiting the move
        keys2 = np.sort(gen.gaussian_keys(keys.size, kd, seed=100 + trial) // kd(2 + trial % 3))
        leaves2, counts2 = port.compute_octree(keys2, 16 + 5 * trial)
        b2, rc2 = assignment(nranks, leaves2, counts2, old=b)
        rb2, rrc2 = assignment(nranks, leaves2, counts2, old=b, lib=ref.lib, prefix="cstone_ref_")
        assert np.array_equal(b2, rb2) and np.array_equal(rc2, rrc2)
        assert np.all(b2[1:] >= b2[:-1])
        # the limiter on its own, with a random previous assignment
        old = np.sort(rng.choice(leaves2, nranks + 1)).astype(kd)
        old[0], old[-1] = leaves2[0], leaves2[-1]
        nb, nrc = limit_shifts(nranks, old, rb, rrc, leaves2, counts2)
        rnb, rnrc = limit_shifts(nranks, old, rb, rrc, leaves2, counts2, lib=ref.lib, prefix="cstone_ref_")
        assert np.array_equal(nb, rnb) and np.array_equal(nrc, rnrc)


@pytest.mark.parametrize("fd", [np.float32, np.float64])
def test_limit_box_shrinking_vs_reference(ref, fd):
    ts = "f32" if fd == np.float32 else "f64"
    lib = _lib()
    rng = np.random.default_rng(3)
    for _ in range(200):
        lo = rng.uniform(-3, 0, 3)
        hi = lo + rng.uniform(0.1, 5, 3)
        prev = np.stack([lo, hi], axis=1).reshape(6).astype(fd)
        shrink = rng.uniform(-0.2, 0.3, 6) * np.repeat(hi - lo, 2) * np.array([1, -1] * 3)
        fit = (prev + shrink).astype(fd)
        out, rout = np.zeros(6, fd), np.zeros(6, fd)
        getattr(lib, "cstone_b200_host_limit_box_shrinking_" + ts)(_p(fit), _p(prev), _p(out))
        getattr(ref.lib, "cstone_ref_limit_box_shrinking_" + ts)(_p(fit), _p(prev), _p(rout))
        assert np.array_equal(out, rout)
        assert np.all(out[0::2] <= fit[0::2]) and np.all(out[1::2] >= fit[1::2])


def test_limit_box_shrinking_golden():
    """[ref: test/unit/sfc/box.cpp limitBoxShrinking] a box that shrank by more than 5 % per side is held back"""
    lib = _lib()
    prev = np.array([0, 1, 0, 1, 0, 1], dtype=np.float64)
    fit = np.array([0.1, 0.9, 0.01, 0.99, -0.5, 2.0], dtype=np.float64)
    out = np.zeros(6)
    lib.cstone_b200_host_limit_box_shrinking_f64(_p(fit), _p(prev), _p(out))
    assert out.tolist() == [0.05, 0.95, 0.01, 0.99, -0.5, 2.0]
    # [ref: test/unit/sfc/box.cpp:175-190 limitBox.belowShrinkLimit]
    prev = np.array([1, 2, 10, 20, 100, 200], dtype=np.float64)
    fit = np.array([1.1, 1.9, 11, 19, 110, 190], dtype=np.float64)
    lib.cstone_b200_host_limit_box_shrinking_f64(_p(fit), _p(prev), _p(out))
    assert np.allclose(out, [1.05, 1.95, 10.5, 19.5, 105, 195], rtol=0, atol=1e-6)
    # [ref: test/unit/sfc/box.cpp:149-159 limitBox.noShrink] a grown box is returned unchanged
    prev = np.array([0, 1, 2, 3, 4, 5], dtype=np.float64)
    fit = np.array([-1, 2, -2, 3, -4, 6], dtype=np.float64)
    lib.cstone_b200_host_limit_box_shrinking_f64(_p(fit), _p(prev), _p(out))
    assert out.tolist() == fit.tolist()


@pytest.mark.parametrize("kd,fd", [(np.uint64, np.float64), (np.uint32, np.float64), (np.uint64, np.float32)])
def test_find_peers_vs_reference(port, kd, fd):
    """findPeersMac [ref: traversal/peers.hpp:62-117]: the host dual traversal of the library (csrc/let_tree.cuh) against
    the reference's, on global trees of clustered clouds: many rank counts, open and periodic boxes, min-MAC and
    vector-MAC opening parameters"""
    import ctypes as C
    import gen
    from oracle import loader
    if not loader.refdom_available():
        pytest.skip("oracle/_ref/liboracle_refdom.so not built")
    from cornerstone_octree_b200 import _lib, ops
    rlib = C.CDLL(loader._REFDOM_PATH)
    sfx = "_%s_%s" % ("k64" if kd == np.uint64 else "k32", "f64" if fd == np.float64 else "f32")
    mine = getattr(_lib.lib, "cstone_b200_host_find_peers" + sfx)
    refp = getattr(rlib, "cstone_refdom_find_peers" + sfx)
    mine.restype = refp.restype = C.c_int
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)
    checked = with_non_peers = 0
    for trial, (nranks, bucket, theta) in enumerate([(2, 64, 0.5), (8, 32, 1.0), (13, 8, 0.4), (64, 4, 0.5)]):
        keys = np.sort(gen.gaussian_keys(12000 + 1000 * trial, kd, seed=40 + trial))
        leaves, counts = port.compute_octree(keys, bucket)
        bnd, _ = assignment(nranks, leaves, counts)
        bnd = np.ascontiguousarray(bnd, dtype=kd)
        for pbc in ((0, 0, 0), (1, 0, 1)):
            lim = np.array([-1.0, 1.0, -1.0, 1.0, -1.0, 1.0], dtype=fd)
            import torch
            box = ops.make_box((-1.0, 1.0), tuple(1 if p else 0 for p in pbc), dtype=torch.float64 if fd == np.float64 else torch.float32)
            pb = np.array(pbc, dtype=np.int32)
            for inv_theta in (np.float32(1.0 / theta + 0.5), np.float32(1.0 / theta + np.sqrt(np.float32(3.0))), np.float32(1.0)):
                for rank in sorted({0, nranks // 2, nranks - 1}):
                    a = np.zeros(nranks, dtype=np.int32)
                    b = np.zeros(nranks, dtype=np.int32)
                    st = mine(1, rank, nranks, ptr(bnd), ptr(leaves), C.c_int32(leaves.size - 1), C.byref(box),
                              C.c_float(float(inv_theta)), ptr(a))
                    assert st == 0
                    st = refp(rank, nranks, ptr(bnd), ptr(leaves), leaves.size - 1, ptr(lim), ptr(pb), C.c_float(float(inv_theta)),
                              ptr(b))
                    assert st == 0
                    assert np.array_equal(a, b), "peers of rank %d of %d (pbc %s, 1/theta %g)" % (rank, nranks, pbc, inv_theta)
                    assert a[rank] == 0
                    checked += 1
                    with_non_peers += int(a.sum() < nranks - 1)
    assert checked >= 60 and with_non_peers > 10


@pytest.mark.parametrize("kd", [np.uint64, np.uint32])
def test_update_treelet_vs_reference(port, kd):
    """the rebalance of a leaf range that is handed to a neighbour rank (focusTransfer -> updateTreelet
    [ref: focus/exchange_focus.hpp:380-405, tree/csarray.hpp:477-488]): merges whose sibling group reaches outside of the
    range, multi-level splits and plain copies, against the reference's function on the same full leaf array"""
    import ctypes as C
    import gen
    from oracle import loader
    if not loader.refdom_available():
        pytest.skip("oracle/_ref/liboracle_refdom.so not built")
    from cornerstone_octree_b200 import _lib
    rlib = C.CDLL(loader._REFDOM_PATH)
    ks = "k64" if kd == np.uint64 else "k32"
    mine = getattr(_lib.lib, "cstone_b200_host_update_treelet_" + ks)
    refp = getattr(rlib, "cstone_refdom_update_treelet_%s_f64" % ks)
    mine.restype = refp.restype = C.c_int
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)
    rng = np.random.default_rng(3)
    cases = changed = 0
    for trial in range(6):
        keys = np.sort(gen.gaussian_keys(20000 + 3000 * trial, kd, seed=60 + trial))
        leaves, counts = port.compute_octree(keys, 8 + 4 * trial)
        # counts of a LATER particle set on the same leaves: merges and (multi-level) splits become due
        keys2 = np.sort(gen.gaussian_keys(keys.size * (1 + trial % 3), kd, seed=160 + trial) // kd(1 + trial % 2))
        counts2 = port.compute_node_counts(leaves, keys2)
        L = leaves.size - 1
        for _ in range(12):
            first = int(rng.integers(0, L - 1))
            last = int(min(L, first + rng.integers(1, max(2, L // 3))))
            for bucket in (4, 8 + 4 * trial, 64):
                cap = 4096 * (last - first) + 1
                a = np.zeros(cap, dtype=kd)
                b = np.zeros(cap, dtype=kd)
                na = mine(ptr(leaves), ptr(counts2), first, last, C.c_uint32(bucket), ptr(a), cap)
                nb = refp(ptr(leaves), ptr(counts2), first, last, C.c_uint(bucket), ptr(b))
                assert na == nb and na > 0, "treelet size, leaves [%d, %d) bucket %d" % (first, last, bucket)
                assert np.array_equal(a[:na], b[:nb])
                cases += 1
                changed += int(na != last - first + 1)
    assert cases >= 200 and changed > 50
