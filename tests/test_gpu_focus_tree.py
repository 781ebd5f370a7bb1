"""The device-resident single-rank focus tree (cstone_b200_focus_tree_*) step by step against the reference's
FocusedOctreeSingleNode compiled from its own headers (oracle/_ref), on the scenario of the reference's unit test
[ref: test/unit/focus/octree_focus.cpp:339-421 computeEssentialTree] and on a particle cloud.  Bit-exact."""
import numpy as np
import pytest
import torch

from oracle import loader
import gpu_adapter as ga
from cornerstone_octree_b200.focus_tree import FocusedOctreeSingleNode

pytestmark = pytest.mark.gpu


def _keys_np(t, kd):
    return t.cpu().numpy().view(np.dtype(kd))


def _step_both(ref_tree, gpu_tree, rbox, gbox, keys, dkeys, fs, fe, mandatory=()):
    want_conv = ref_tree.update(rbox, keys, fs, fe, mandatory)
    got_conv = gpu_tree.update(gbox, dkeys, fs, fe, mandatory)
    wl, wc = ref_tree.leaves_and_counts()
    np.testing.assert_array_equal(_keys_np(gpu_tree.tree_leaves(), keys.dtype), wl)
    np.testing.assert_array_equal(gpu_tree.leaf_counts().cpu().numpy().view(np.uint32), wc)
    assert got_conv == want_conv
    return want_conv


def _octant_node_count(tree, kd):
    ml = loader.max_level(kd)
    rng = 1 << (3 * (ml - 1))
    t = tree.astype(object)
    counts = []
    for octant in range(8):
        a, b = octant * rng, octant * rng + rng
        ia, ib = np.searchsorted(tree, np.array(a, dtype=tree.dtype)), np.searchsorted(tree, np.array(b, dtype=tree.dtype))
        assert int(t[ia]) == a and int(t[ib]) == b
        counts.append(int(ib - ia))
    return sorted(counts)[:7]


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
def test_compute_essential_tree(ref, kd):
    n = 200000
    ml = loader.max_level(kd)
    top = 1 << (3 * ml)
    keys = (np.arange(n, dtype=np.float64) * (float(top) / n)).astype(kd)
    rbox = loader.make_box((-1, 1), dtype=np.float64)
    gbox = ga.conv_box(rbox)
    dkeys = ga.to_dev(keys)
    td = torch.int32 if kd == np.uint32 else torch.int64
    rt = loader.RefFocusTree(ref, 16, 0.9, kd)
    gt = FocusedOctreeSingleNode(16, 0.9, key_dtype=td)
    ref_counts = [29, 302, 302, 302, 1184, 1184, 1184]  # the reference test's known answer

    def pad(prefix, length):
        return prefix << (3 * ml - length)

    fs, fe = 1, pad(1, 3)
    _step_both(rt, gt, rbox, gbox, keys, dkeys, fs, fe)
    # worst case: a focus boundary on the deepest level must be resolved after ONE step (forced injection)
    assert int(_keys_np(gt.tree_leaves(), kd)[1]) == fs
    steps = 0
    while not _step_both(rt, gt, rbox, gbox, keys, dkeys, fs, fe):
        steps += 1
        assert steps < 40
    if kd is not None:  # both key widths give the reference test's known answer
        assert _octant_node_count(_keys_np(gt.tree_leaves(), kd), kd) == ref_counts

    fs, fe = pad(7, 3), top - 1
    while not _step_both(rt, gt, rbox, gbox, keys, dkeys, fs, fe):
        steps += 1
        assert steps < 80
    fs, fe = 0, pad(1, 3)
    while not _step_both(rt, gt, rbox, gbox, keys, dkeys, fs, fe):
        steps += 1
        assert steps < 120
    if kd is not None:  # both key widths give the reference test's known answer
        assert _octant_node_count(_keys_np(gt.tree_leaves(), kd), kd) == ref_counts
    # node arrays are consistent: root count = all particles, marks only outside of nothing-but-the-focus
    prefixes, child_offsets, counts, macs = gt.node_arrays()
    assert int(counts[0].item()) == n and int(macs.sum().item()) > 0
    rt.close()
    gt.close()


@pytest.mark.parametrize("fd", [np.float32, np.float64])
@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
def test_focus_tree_cloud_with_mandatory_keys(ref, kd, fd):
    n = 100000
    rng = np.random.default_rng(31)
    rbox = loader.make_box((-1, 1), (loader.PERIODIC, loader.OPEN, loader.OPEN), dtype=fd)
    gbox = ga.conv_box(rbox)
    xyz = np.clip(rng.normal(0, 0.25, size=(3, n)), -0.999, 0.999).astype(fd)
    keys = np.sort(ref.compute_sfc_keys(loader.HILBERT, xyz[0].copy(), xyz[1].copy(), xyz[2].copy(), rbox, kdtype=kd))
    dkeys = ga.to_dev(keys)
    td = torch.int32 if kd == np.uint32 else torch.int64
    rt = loader.RefFocusTree(ref, 32, 0.7, kd)
    gt = FocusedOctreeSingleNode(32, 0.7, key_dtype=td)
    top = 1 << (3 * loader.max_level(kd))
    # focus = an SFC slice as a domain assignment would give; mandatory keys = deep, unsorted, with duplicates
    fs, fe = int(keys[n // 3]), int(keys[2 * n // 3])
    mand = [int(keys[n // 10]), int(keys[9 * n // 10]) | 1, int(keys[n // 10]), top // 8 + 5]
    steps = 0
    while not _step_both(rt, gt, rbox, gbox, keys, dkeys, fs, fe, mand):
        steps += 1
        assert steps < 60
    leaves = _keys_np(gt.tree_leaves(), kd)
    for k in [fs, fe] + mand:
        assert leaves[np.searchsorted(leaves, np.array(k, dtype=kd))] == k
    # moving focus: the tree follows
    fs, fe = int(keys[n // 2]), int(keys[5 * n // 6])
    while not _step_both(rt, gt, rbox, gbox, keys, dkeys, fs, fe, mand[:1]):
        steps += 1
        assert steps < 120
    rt.close()
    gt.close()


def _mac_refinement_tree(kd):
    """OctreeMaker{}.divide().divide(0).divide(0,j) for j in 0..7 .divide(0,0,0): level 3 in the first octant (level 4
    in its first cell), level 1 elsewhere  [ref: test/unit/focus/octree_focus.cpp:436-447]"""
    ml = loader.max_level(kd)
    r = [1 << (3 * (ml - l)) for l in range(6)]
    keys = set(i * r[1] for i in range(9))
    keys |= set(i * r[2] for i in range(8))
    keys |= set(i * r[3] for i in range(64))
    keys |= set(i * r[4] for i in range(8))
    return np.array(sorted(keys), dtype=kd)


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
@pytest.mark.parametrize("fd", [np.float32, np.float64])
def test_mac_refinement(ref, kd, fd):
    ml = loader.max_level(kd)
    leaves0 = _mac_refinement_tree(kd)
    assert leaves0.size - 1 == 78
    rbox = loader.make_box((0, 1), dtype=fd)
    gbox = ga.conv_box(rbox)
    inv_theta = float(np.float32(np.sqrt(3) / 2 + 1e-6))
    td = torch.int32 if kd == np.uint32 else torch.int64
    oct1 = 1 << (3 * (ml - 1))  # decodePlaceholderBit(011): end of the first octant
    o0101 = 1 << (3 * (ml - 2))  # decodePlaceholderBit(0101)
    o0107 = 7 << (3 * (ml - 2))  # decodePlaceholderBit(0107)
    cases = {"fullSurface": (oct1, oct1, 64 + 7 + 3 * (4 + 4 * 8) + 3 * (6 + 2 * 8) + (7 + 8)),
             "noSurface": (o0101, oct1, 78),
             "partialSurface": (0, o0107, 78 + 5 * 7)}
    for name, (old_fs, old_fe, want_leaves) in cases.items():
        gt = FocusedOctreeSingleNode(16, 0.5, key_dtype=td)
        gt.set_leaves(ga.to_dev(leaves0))
        leaves = leaves0
        steps = 0
        while True:
            leaves, want_conv = ref.mac_refine_step(leaves, old_fs, old_fe, 0, oct1, inv_theta, rbox)
            got_conv = gt.mac_refine(gbox, old_fs, old_fe, 0, oct1, inv_theta)
            np.testing.assert_array_equal(_keys_np(gt.tree_leaves(), kd), leaves, err_msg=name)
            assert got_conv == want_conv
            steps += 1
            assert steps < 20
            if want_conv:
                break
        if fd == np.float64:
            assert leaves.size - 1 == want_leaves, name  # the reference test's known answers (double)
        # unchanged focus: nothing to do
        assert gt.mac_refine(gbox, 0, oct1, 0, oct1, inv_theta)
        gt.close()
