"""Pin the port oracle (oracle/cstone_oracle.cpp) against the reference's own CPU code compiled from
/root/reference/include (oracle/_ref).  Bit-exact on every stage of the hot path."""
import numpy as np
import pytest

from oracle import loader
import gen
import pipeline

CASES = [
    # name, generator, n, kdtype, fdtype, kind, bucket, boundaries
    ("uniform-k64-f64-hilbert", "uniform", 20000, np.uint64, np.float64, 1, 64, (0, 0, 0)),
    ("uniform-k64-f64-hilbert-pbc", "uniform", 20000, np.uint64, np.float64, 1, 64, (1, 1, 1)),
    ("uniform-k32-f32-hilbert-pbc", "uniform", 15000, np.uint32, np.float32, 1, 16, (1, 0, 1)),
    ("gaussian-k64-f32-hilbert", "gaussian", 20000, np.uint64, np.float32, 1, 16, (0, 0, 0)),
    ("plummer-k64-f64-hilbert", "plummer", 20000, np.uint64, np.float64, 1, 16, (0, 0, 0)),
    ("gaussian-k32-f64-hilbert", "gaussian", 20000, np.uint32, np.float64, 1, 8, (0, 1, 0)),
]


def make_case(genname, n, fdtype, boundaries):
    if genname == "uniform":
        lim = (0.0, 1.0)
        x, y, z = gen.uniform(n, lim, fdtype)
    elif genname == "gaussian":
        lim = (-1.0, 1.0)
        x, y, z = gen.gaussian(n, lim, fdtype)
    else:
        (x, y, z), lim = gen.plummer(n, fdtype)
    box = loader.make_box(lim, boundaries, fdtype)
    rng = np.random.default_rng(7)
    ext = float(box.lengths[0])
    h = (0.5 * ext * np.cbrt(3 * 60 / (4 * np.pi * n)) * (0.8 + 0.4 * rng.random(n))).astype(fdtype)
    return x, y, z, h, box


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_port_matches_reference(case, port, ref):
    _, genname, n, kd, fd, kind, bucket, bnd = case
    x, y, z, h, box = make_case(genname, n, fd, bnd)
    ngmax = 150
    a = pipeline.run(port, kind, x, y, z, h, box, kd, bucket, ngmax)
    b = pipeline.run(ref, kind, x, y, z, h, box, kd, bucket, ngmax)
    assert pipeline.compare(a, b, ngmax) == []
    assert a["flags"].sum() > 0
    assert a["neighbors_count"].mean() > 5


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
@pytest.mark.parametrize("fd", [np.float32, np.float64])
def test_morton_keys_match_reference(kd, fd, port, ref):
    x, y, z = gen.uniform(50000, (-1, 1), fd)
    box = loader.make_box((-1, 1), dtype=fd)
    assert np.array_equal(port.compute_sfc_keys(0, x, y, z, box, kd), ref.compute_sfc_keys(0, x, y, z, box, kd))


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
def test_update_steps_match_reference(kd, port, ref):
    """tree and counts after EVERY update step, not only at convergence"""
    keys = gen.gaussian_keys(30000, kd)
    root = np.array([0, 2 ** (3 * loader.max_level(kd))], dtype=kd)
    ta, ca = root.copy(), np.array([keys.size], dtype=np.uint32)
    tb, cb = ta.copy(), ca.copy()
    for _ in range(12):
        ta, ca, conv_a = port.update_octree(keys, 16, ta, ca)
        tb, cb, conv_b = ref.update_octree(keys, 16, tb, cb)
        assert np.array_equal(ta, tb) and np.array_equal(ca, cb) and conv_a == conv_b
        if conv_a:
            break
    assert conv_a
    # ops on the converged tree for a different bucket size incl. merges
    for bucket in (4, 64, 1000):
        oa, _ = port.rebalance_decision(ta, ca, bucket)
        ob, _ = ref.rebalance_decision(tb, cb, bucket)
        assert np.array_equal(oa, ob)
        na, sa = port.rebalance_tree(ta, oa)
        nb, sb = ref.rebalance_tree(tb, ob)
        assert np.array_equal(na, nb) and np.array_equal(sa, sb)


def test_removed_keys_untouched(port, ref):
    x, y, z = gen.uniform(1000)
    box = loader.make_box((0, 1))
    for kd in (np.uint32, np.uint64):
        keys = np.zeros(1000, dtype=kd)
        keys[::7] = 2 ** (3 * loader.max_level(kd))
        a = port.compute_sfc_keys(1, x, y, z, box, kd, keys.copy())
        b = ref.compute_sfc_keys(1, x, y, z, box, kd, keys.copy())
        assert np.array_equal(a, b) and (a[::7] == keys[::7]).all()
