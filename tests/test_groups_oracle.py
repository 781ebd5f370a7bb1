"""oracle/groups_np.py against the reference's own known-answer vectors [ref: test/unit_cuda/traversal/groups.cu]."""
import numpy as np

from oracle import groups_np as G


def test_fixed_groups_t0():
    # TEST(TargetGroups, t0)
    assert G.fixed_groups(4, 34, 8).tolist() == [4, 12, 20, 28, 34]


def test_make_splits_vectors():
    # TEST(TargetGroups, makeSplits), masks given as (segment 0, segment 1)
    assert G.make_splits((0, 0)) == [64]
    assert G.make_splits((1, 0)) == [1, 63]
    assert G.make_splits((0, 1 << 30)) == [63, 1]
    assert G.make_splits((2, 0)) == [2, 62]
    assert G.make_splits((3, 0)) == [1, 1, 62]
    assert G.make_splits((1 << 31, 1)) == [32, 1, 31]
    assert G.make_splits((0, 8)) == [36, 28]
    l = G.make_splits((0xFFFFFFFF, 0x6FFFFFFF))
    assert [l[i] for i in range(63)] == [2 if i == 60 else 1 for i in range(63)]
    l = G.make_splits((0xFFFFFFFF, 0x7FFFFFFF))
    assert l[:63] == [1] * 63


def group_volumes_case():
    """inputs of TEST(TargetGroups, groupVolumes): OctreeMaker{}.divide().divide(2) leaves, counts, x=y=z=iota"""
    first, last = 4, 128
    r1 = 1 << 60  # nodeRange<uint64>(1)
    r2 = 1 << 57
    leaves = [0, r1, 2 * r1] + [2 * r1 + i * r2 for i in range(1, 8)] + [i * r1 for i in range(3, 9)]
    leaves = np.array(leaves, dtype=np.uint64)
    counts = [4, 1, 8, 8, 8, 8, 31, 8, 8, 8, 16, 16, 16, 0, 0]
    layout = np.concatenate(([0], np.cumsum(counts))).astype(np.uint32)
    x = np.arange(last, dtype=np.float64)
    x[5] -= 0.01
    inv = [np.float64(1.0) / np.float64(last)] * 3
    dist_crit = np.cbrt(float(last) ** 3 / 64)
    return first, last, leaves, layout, x, inv, dist_crit


def test_group_volumes_vectors():
    first, last, leaves, layout, x, inv, dist_crit = group_volumes_case()
    assert leaves.size == 16
    tol = np.float32(np.sqrt(3.0) / dist_crit * 1.01)
    _, ns = G.group_split_masks(first, last, x, x, x, leaves, layout, inv, 64, tol)
    assert ns.tolist() == [2, 1]
    tol2 = np.float32(np.sqrt(3.0) / dist_crit * 0.99)
    _, ns = G.group_split_masks(first, last, x, x, x, leaves, layout, inv, 64, tol2)
    assert ns.tolist() == [64, 60]
    assert G.group_splits(first, last, x, x, x, leaves, layout, inv, 64, tol).tolist() == [4, 6, 68, 128]
