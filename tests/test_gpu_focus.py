"""Parity of the focus-tree update kernels (rebalanceDecisionEssentialGpu, macRefineDecisionGpu,
protectAncestorsGpu, enforceKeysGpu, rangeCountGpu, geoMacSpheresGpu, setMacGpu, markMacsGpu) with the reference's
CPU functions compiled from its own headers (oracle/_ref).  Integer outputs bit-exact; the MAC spheres are evaluated
with the reference's expression order under -fmad=false and compared bit-exact too."""
import numpy as np
import pytest
import torch

from oracle import loader
import gpu_adapter as ga
from cornerstone_octree_b200 import ops

pytestmark = pytest.mark.gpu


def _cloud(n, seed, fdtype):
    rng = np.random.default_rng(seed)
    xyz = np.clip(rng.normal(0.0, 0.18, size=(3, n)), -0.999, 0.999).astype(fdtype)
    return xyz[0].copy(), xyz[1].copy(), xyz[2].copy()


def _tree(ref, kd, fd, n=60000, bucket=16, seed=7, boundaries=(loader.OPEN,) * 3):
    box = loader.make_box((-1, 1), boundaries, dtype=fd)
    x, y, z = _cloud(n, seed, fd)
    keys = np.sort(ref.compute_sfc_keys(loader.HILBERT, x, y, z, box, kdtype=kd))
    tree, counts = ref.compute_octree(keys, bucket)
    oc = ref.build_octree(tree)
    node_counts = np.zeros(oc["num_nodes"], dtype=np.uint32)
    node_counts[oc["leaf_to_internal"][oc["num_internal_nodes"]:]] = counts
    node_counts = ref.upsweep_sum(oc["level_range"], oc["child_offsets"], node_counts)
    return box, keys, tree, counts, oc, node_counts


KD = [np.uint32, np.uint64]


@pytest.mark.parametrize("kd", KD)
def test_rebalance_decision_essential_and_protect(ref, kd):
    box, keys, tree, counts, oc, nc = _tree(ref, kd, np.float64)
    rng = np.random.default_rng(3)
    d = ga.oct_to_dev(oc)
    L = oc["num_leaf_nodes"]
    for trial, bucket in enumerate((16, 64, 4)):
        macs = (rng.random(oc["num_nodes"]) < (0.3 + 0.2 * trial)).astype(np.int8)
        a, b = sorted(rng.integers(0, L + 1, size=2))
        want = ref.rebalance_decision_essential(oc, nc, macs, tree[a], tree[b], bucket)
        got = ops.rebalance_decision_essential(d, ga.to_dev(nc), ga.to_dev(macs), int(tree[a]), int(tree[b]), bucket)
        np.testing.assert_array_equal(got.cpu().numpy(), want)

        want_p, want_conv = ref.protect_ancestors(oc, want)
        conv = ops.protect_ancestors(d, got)
        np.testing.assert_array_equal(got.cpu().numpy(), want_p)
        assert conv == want_conv
    # the converged case: every op already 1
    ones = torch.ones(oc["num_nodes"], dtype=torch.int32, device="cuda")
    assert ops.protect_ancestors(d, ones) == ref.protect_ancestors(oc, np.ones(oc["num_nodes"], np.int32))[1]


@pytest.mark.parametrize("kd", KD)
def test_mac_refine_decision(ref, kd):
    box, keys, tree, counts, oc, nc = _tree(ref, kd, np.float64, n=20000)
    rng = np.random.default_rng(5)
    macs = (rng.random(oc["num_nodes"]) < 0.5).astype(np.int8)
    L = oc["num_leaf_nodes"]
    d = ga.oct_to_dev(oc)
    for a, b in ((0, L), (0, 0), (L // 3, 2 * L // 3), (L - 1, L)):
        want = ref.mac_refine_decision(oc, macs, a, b)
        got = ops.mac_refine_decision(d, ga.to_dev(macs), a, b)
        np.testing.assert_array_equal(got.cpu().numpy(), want)


@pytest.mark.parametrize("kd", KD)
def test_enforce_keys(ref, kd):
    box, keys, tree, counts, oc, nc = _tree(ref, kd, np.float64, n=30000)
    rng = np.random.default_rng(11)
    d = ga.oct_to_dev(oc)
    L = oc["num_leaf_nodes"]
    bits = 8 * np.dtype(kd).itemsize
    top = 1 << (3 * loader.max_level(kd))

    def forced(kind):
        if kind == "present":  # keys of the tree itself: nothing to do
            return tree[rng.integers(0, L + 1, size=40)].copy()
        if kind == "one_deeper":  # first-child boundaries one level below a leaf
            i = rng.integers(0, L, size=25)
            w = (tree[i + 1] - tree[i])
            ok = w >= 8
            return (tree[i][ok] + (w[ok] // 8) * rng.integers(1, 8, size=ok.sum()).astype(kd)).astype(kd)
        if kind == "deep":  # arbitrary keys, several levels below the leaves
            return np.sort(rng.integers(0, top, size=30, dtype=np.uint64)).astype(kd)
        return np.array([0, top], dtype=kd)

    for kind in ("present", "one_deeper", "deep", "ends"):
        fk = forced(kind)
        macs = (rng.random(oc["num_nodes"]) < 0.5).astype(np.int8)
        a, b = sorted(rng.integers(0, L + 1, size=2))
        ops0 = ref.rebalance_decision_essential(oc, nc, macs, tree[a], tree[b], 16)
        ops0, _ = ref.protect_ancestors(oc, ops0)
        want, want_status = ref.enforce_keys(fk, oc, ops0)
        got = ga.to_dev(ops0)
        status = ops.enforce_keys(ga.to_dev(fk), d, got)
        assert status == want_status, kind
        np.testing.assert_array_equal(got.cpu().numpy(), want, err_msg=kind)
    assert bits in (32, 64)


@pytest.mark.parametrize("kd", KD)
def test_range_count(ref, kd):
    box, keys, tree, counts, oc, nc = _tree(ref, kd, np.float64, n=50000, bucket=8)
    # a coarser focus tree over the same keys: every focus leaf is a union of global leaves
    ftree, _ = ref.compute_octree(keys, 64)
    nf = ftree.size - 1
    rng = np.random.default_rng(2)
    idx = np.sort(rng.choice(nf, size=nf // 2, replace=False)).astype(np.int32)
    big = counts.copy()
    big[::7] = 0xF0000000  # saturation of the sum at 2^32 - 1
    for c in (counts, big):
        want = ref.range_count(tree, c, ftree, idx)
        got = ops.range_count(ga.to_dev(tree), ga.to_dev(c), ga.to_dev(ftree), ga.to_dev(idx))
        np.testing.assert_array_equal(got.cpu().numpy().view(np.uint32), want)
    # empty index list
    got = ops.range_count(ga.to_dev(tree), ga.to_dev(counts), ga.to_dev(ftree), ga.to_dev(idx[:0]))
    assert int(got.abs().sum()) == 0


@pytest.mark.parametrize("fd", [np.float32, np.float64])
@pytest.mark.parametrize("kd", KD)
@pytest.mark.parametrize("pbc", [False, True])
def test_mac_spheres_and_mark_macs(ref, kd, fd, pbc):
    bnd = (loader.PERIODIC,) * 3 if pbc else (loader.OPEN,) * 3
    box, keys, tree, counts, oc, nc = _tree(ref, kd, fd, n=40000, bucket=16, boundaries=bnd)
    gbox = ga.conv_box(box)
    d = ga.oct_to_dev(oc)
    td = torch.float32 if fd == np.float32 else torch.float64
    L = oc["num_leaf_nodes"]
    for inv_theta in (2.0, 1.0 / 0.6):
        want_geo = ref.geo_mac_spheres(oc["prefixes"], inv_theta, box, fdtype=fd)
        got_geo = ops.geo_mac_spheres(d.prefixes, inv_theta, gbox, dtype=td)
        np.testing.assert_array_equal(got_geo.cpu().numpy(), want_geo)

        # expansion centres displaced from the geometric ones, some nodes flagged empty (mac 0)
        rng = np.random.default_rng(13)
        cen = want_geo.copy()
        cen[:, :3] += (rng.uniform(-0.01, 0.01, size=(cen.shape[0], 3))).astype(fd)
        cen[:, 3] = 1
        cen[rng.random(cen.shape[0]) < 0.1, 3] = 0
        cen[:73, 3] = 1  # keep the top levels open, or the traversal ends at the root
        want_vec = ref.set_mac(oc["prefixes"], cen, inv_theta, box)
        got_vec = ops.set_mac(d.prefixes, ga.to_dev(cen), inv_theta, gbox)
        np.testing.assert_array_equal(got_vec.cpu().numpy(), want_vec)

        for a, b in ((L // 4, L // 2), (0, L // 10), (L - 3, L)):
            focus = tree[a:b + 1].copy()
            for limit in (False, True):
                for c4 in (want_geo, want_vec):
                    want = ref.mark_macs(oc, c4, box, focus, limit)
                    got = ops.mark_macs(d, ga.to_dev(c4), gbox, ga.to_dev(focus), limit)
                    np.testing.assert_array_equal(got.cpu().numpy(), want)
                    assert want.sum() > 0


@pytest.mark.parametrize("fd", [np.float32, np.float64])
def test_move_centers(fd):
    rng = np.random.default_rng(0)
    src = rng.normal(size=(1000, 3)).astype(fd)
    got = ops.move_centers(ga.to_dev(src)).cpu().numpy()
    np.testing.assert_array_equal(got[:, :3], src)
    np.testing.assert_array_equal(got[:, 3], np.ones(1000, dtype=fd))


@pytest.mark.parametrize("types", [(np.float64, np.float64, np.float64), (np.float64, np.float32, np.float64),
                                   (np.float32, np.float32, np.float32)])
@pytest.mark.parametrize("kd", KD)
def test_source_centers(ref, kd, types):
    """computeLeafSourceCenterGpu + upsweepCentersGpu: sums in the reference's order, bit-identical"""
    tc, tm, tf = types
    n = 50000
    box = loader.make_box((-1, 1), dtype=tc)
    x, y, z = _cloud(n, 9, tc)
    keys = ref.compute_sfc_keys(loader.HILBERT, x, y, z, box, kdtype=kd)
    order = np.argsort(keys, kind="stable")
    keys, x, y, z = keys[order], x[order].copy(), y[order].copy(), z[order].copy()
    rng = np.random.default_rng(4)
    m = rng.uniform(0.1, 2.0, size=n).astype(tm)
    m[::17] *= -1  # addBody weighs with |m|
    tree, counts = ref.compute_octree(keys, 32)
    oc = ref.build_octree(tree)
    layout = np.concatenate(([0], np.cumsum(counts, dtype=np.uint64))).astype(np.uint32)
    d = ga.oct_to_dev(oc)

    want_leaf = ref.compute_leaf_source_center(x, y, z, m, oc, layout, fdtype=tf)
    got = ops.compute_leaf_source_center(ga.to_dev(x), ga.to_dev(y), ga.to_dev(z), ga.to_dev(m), d, ga.to_dev(layout))
    np.testing.assert_array_equal(got.cpu().numpy(), want_leaf)

    want = ref.upsweep_centers(oc["level_range"], oc["child_offsets"], want_leaf)
    ops.upsweep_centers(oc["level_range"], d.child_offsets, got)
    np.testing.assert_array_equal(got.cpu().numpy(), want)
    # the root holds the total |mass| up to summation order and a centre inside the cloud
    assert abs(want[0, 3] - np.abs(m.astype(np.float64)).sum()) < 1e-3 * n
    assert np.all(np.abs(want[0, :3]) < 0.1)


class _GpuFocusImpl:
    """focus_kat interface over the C ABI"""

    def enforce(self, keys, oc, node_ops):
        d = ga.oct_to_dev(oc)
        got = ga.to_dev(node_ops)
        status = ops.enforce_keys(ga.to_dev(keys), d, got)
        return got.cpu().numpy(), status

    def protect(self, oc, node_ops):
        d = ga.oct_to_dev(oc)
        got = ga.to_dev(node_ops)
        conv = ops.protect_ancestors(d, got)
        return got.cpu().numpy(), conv

    def decision(self, oc, counts, macs, fs, fe, bucket):
        d = ga.oct_to_dev(oc)
        return ops.rebalance_decision_essential(d, ga.to_dev(counts), ga.to_dev(macs), fs, fe, bucket).cpu().numpy()


def test_reference_known_answers(port):
    """the reference unit tests' vectors for enforceKeys / protectAncestors, through the CUDA library"""
    import focus_kat
    focus_kat.check_key_enforcement(_GpuFocusImpl(), port.build_octree)
    focus_kat.check_protect_ancestors(_GpuFocusImpl(), port.build_octree)
    focus_kat.check_rebalance_decision(_GpuFocusImpl(), port.build_octree, port.upsweep_sum)
