"""Target groups (computeFixedGroups / computeGroupSplits) through the C ABI against oracle/groups_np.py
and the reference's known-answer vectors [ref: test/unit_cuda/traversal/groups.cu].  Integer outputs, bit-exact."""
import numpy as np
import pytest
import torch

from oracle import groups_np as G
from oracle import loader
import gpu_adapter as ga
from cornerstone_octree_b200 import ops
from test_groups_oracle import group_volumes_case

pytestmark = pytest.mark.gpu


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


def test_fixed_groups():
    assert _u32(ops.compute_fixed_groups(4, 34, 8)).tolist() == [4, 12, 20, 28, 34]
    for first, last, gs in ((0, 1000, 64), (17, 18, 32), (5, 5, 32), (3, 67, 64), (0, 1 << 20, 64)):
        np.testing.assert_array_equal(_u32(ops.compute_fixed_groups(first, last, gs)), G.fixed_groups(first, last, gs))


def test_group_volumes_reference_vectors():
    first, last, leaves, layout, x, inv, dist_crit = group_volumes_case()
    box = ga.conv_box(loader.make_box((0, last), dtype=np.float64))
    dx, dl, dy = ga.to_dev(x), ga.to_dev(leaves), ga.to_dev(layout)
    h = torch.zeros_like(dx)
    tol = float(np.float32(np.sqrt(3.0) / dist_crit * 1.01))
    got = ops.compute_group_splits(first, last, dx, dx, dx, h, dl, dy, box, 64, tol)
    assert _u32(got).tolist() == [4, 6, 68, 128]
    tol = float(np.float32(np.sqrt(3.0) / dist_crit * 0.99))
    got = _u32(ops.compute_group_splits(first, last, dx, dx, dx, h, dl, dy, box, 64, tol))
    assert got.size == 64 + 60 + 1 and got[0] == 4 and got[-1] == 128 and (np.diff(got) == 1).all()


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
@pytest.mark.parametrize("types", [(np.float64, np.float64), (np.float64, np.float32), (np.float32, np.float32)])
@pytest.mark.parametrize("group_size", [32, 64])
def test_group_splits_cloud(port, kd, types, group_size):
    tc, th = types
    n = 20000
    rng = np.random.default_rng(21)
    box = loader.make_box((-1, 1), dtype=tc)
    xyz = np.clip(rng.normal(0, 0.3, size=(3, n)), -0.999, 0.999).astype(tc)
    keys = port.compute_sfc_keys(loader.HILBERT, xyz[0].copy(), xyz[1].copy(), xyz[2].copy(), box, kdtype=kd)
    order = np.argsort(keys, kind="stable")
    keys = keys[order]
    x, y, z = (np.ascontiguousarray(a[order]) for a in xyz)
    tree, counts = port.compute_octree(keys, 16)
    layout = np.concatenate(([0], np.cumsum(counts, dtype=np.uint64))).astype(np.uint32)
    inv = [box.inv_lengths[d] for d in range(3)]
    h = np.full(n, 0.01, dtype=th)
    gbox = ga.conv_box(box)
    for first, last in ((0, n), (37, n - 11), (100, 101), (64, 64 + group_size), (9, 9)):
        for tol in (0.9, 2.0, 50.0):
            want = G.group_splits(first, last, x, y, z, tree, layout, inv, group_size, tol, th=th)
            got = ops.compute_group_splits(first, last, ga.to_dev(x), ga.to_dev(y), ga.to_dev(z), ga.to_dev(h),
                                           ga.to_dev(tree), ga.to_dev(layout), gbox, group_size, tol)
            np.testing.assert_array_equal(_u32(got), want, err_msg=str((first, last, tol)))
            if last > first:
                d = np.diff(want.astype(np.int64))
                assert (d > 0).all() and d.max() <= group_size
    # no smoothing lengths given
    got = ops.compute_group_splits(0, n, ga.to_dev(x), ga.to_dev(y), ga.to_dev(z), None, ga.to_dev(tree),
                                   ga.to_dev(layout), gbox, group_size, 2.0)
    want = G.group_splits(0, n, x, y, z, tree, layout, inv, group_size, 2.0, th=tc)
    np.testing.assert_array_equal(_u32(got), want)


def test_group_splits_rejects_other_sizes():
    from cornerstone_octree_b200._lib import CstoneError
    first, last, leaves, layout, x, inv, dist_crit = group_volumes_case()
    box = ga.conv_box(loader.make_box((0, last), dtype=np.float64))
    dx = ga.to_dev(x)
    with pytest.raises(CstoneError):
        ops.compute_group_splits(first, last, dx, dx, dx, None, ga.to_dev(leaves), ga.to_dev(layout), box, 128, 1.0)
