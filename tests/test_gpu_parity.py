"""GPU parity tests proper: the CUDA library (through its C ABI) against the CPU checkers on the same seeded
inputs, bit-exact on every stage of the hot path, plus the reference's golden vectors (kat.py)."""
import numpy as np
import pytest

import gen
import kat
import pipeline
from oracle import loader
from test_oracle_vs_ref import CASES, make_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import gpu_adapter
    return gpu_adapter.GpuImpl()


KATS = [kat.check_morton, kat.check_hilbert, kat.check_sfc_keys_box, kat.check_sort, kat.check_node_counts,
        kat.check_rebalance, kat.check_compute_octree, kat.check_octree, kat.check_find_halos]


@pytest.mark.parametrize("check", KATS, ids=[c.__name__ for c in KATS])
def test_known_answers(check, gpu):
    check(gpu)


def test_neighbors_vs_all2all(gpu):
    kat.check_find_neighbors(gpu, n=1200)


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_pipeline_matches_port(case, gpu, port):
    _, genname, n, kd, fd, kind, bucket, bnd = case
    x, y, z, h, box = make_case(genname, n, fd, bnd)
    ngmax = 150
    a = pipeline.run(port, kind, x, y, z, h, box, kd, bucket, ngmax)
    b = pipeline.run(gpu, kind, x, y, z, h, box, kd, bucket, ngmax)
    assert pipeline.compare(a, b, ngmax) == []


def test_pipeline_matches_reference_build(gpu, ref):
    """against the reference's own code (prebuilt oracle/_ref travels to the GPU box)"""
    x, y, z, h, box = make_case("plummer", 50000, np.float64, (0, 0, 0))
    a = pipeline.run(ref, 1, x, y, z, h, box, np.uint64, 64, 150)
    b = pipeline.run(gpu, 1, x, y, z, h, box, np.uint64, 64, 150)
    assert pipeline.compare(a, b, 150) == []


def test_ngmax_truncation_keeps_dfs_order(gpu, port):
    """count > ngmax: the kept subset is the first ngmax in depth-first order [ref: findneighbors.hpp:138-139]"""
    x, y, z, h, box = make_case("uniform", 20000, np.float64, (1, 1, 1))
    h *= 1.6
    a = pipeline.run(port, 1, x, y, z, h, box, np.uint64, 64, 32)
    b = pipeline.run(gpu, 1, x, y, z, h, box, np.uint64, 64, 32)
    assert (a["neighbors_count"] > 32).mean() > 0.5
    assert pipeline.compare(a, b, 32) == []


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
def test_update_steps_match_port(kd, gpu, port):
    keys = gen.gaussian_keys(200000, kd)
    root = np.array([0, 2 ** (3 * loader.max_level(kd))], dtype=kd)
    ta, ca = root.copy(), np.array([keys.size], dtype=np.uint32)
    tb, cb = ta.copy(), ca.copy()
    for _ in range(12):
        ta, ca, conv_a = port.update_octree(keys, 16, ta, ca)
        tb, cb, conv_b = gpu.update_octree(keys, 16, tb, cb)
        assert np.array_equal(ta, tb) and np.array_equal(ca, cb) and conv_a == conv_b
        if conv_a:
            break
    assert conv_a


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
@pytest.mark.parametrize("n", [0, 1, 31, 4096, 4097, 100003, 3000000])
def test_sort_sizes(kd, n, gpu):
    """ragged / empty sizes; stability; permutation and keys bit-exact against numpy's stable sort"""
    rng = np.random.default_rng(n)
    bits = 30 if kd == np.uint32 else 63
    keys = rng.integers(0, 2 ** bits, n, dtype=np.uint64).astype(kd)
    if n > 10:
        keys[rng.integers(0, n, n // 3)] = keys[0]          # heavy duplicates
        keys[rng.integers(0, n, 5)] = 2 ** bits               # removeKey sorts last
    sk, order = gpu.sort_by_key(keys)
    ref = np.argsort(keys, kind="stable").astype(np.uint32)
    assert np.array_equal(order, ref) and np.array_equal(sk, keys[ref])
    # re-sort with an existing payload (updateMap)
    if n:
        vals = rng.integers(0, 2 ** 31, n).astype(np.uint32)
        sk2, v2 = gpu.sort_by_key(keys, vals)
        assert np.array_equal(sk2, keys[ref]) and np.array_equal(v2, vals[ref])


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
@pytest.mark.parametrize("pattern", ["low16", "low24", "high-only", "middle-constant", "all-equal", "placeholder-prefixes"])
def test_sort_uniform_digit_passes(kd, pattern, gpu):
    """passes whose digit is the same for every key are skipped in pairs or turned into a straight copy
    (radix_sort.cu binOffsetsKernel): same permutation as numpy's stable sort, with and without a payload"""
    n = 300007
    rng = np.random.default_rng(5)
    bits = 30 if kd == np.uint32 else 63
    r = rng.integers(0, 2 ** 62, n, dtype=np.uint64)
    if pattern == "low16":
        keys = r & 0xffff
    elif pattern == "low24":
        keys = r & 0xffffff
    elif pattern == "high-only":
        keys = (r & 0xff) << (bits - 8)
    elif pattern == "middle-constant":
        keys = (r & ((1 << bits) - 1)) & ~np.uint64(0xff00) | np.uint64(0x5a00)
    elif pattern == "all-equal":
        keys = np.full(n, 12345, dtype=np.uint64)
    else:  # Warren-Salmon keys of a depth-5 tree: 16 significant bits and the placeholder bit
        lvl = rng.integers(0, 6, n).astype(np.uint64)
        keys = (np.uint64(1) << (3 * lvl)) | (r & ((np.uint64(1) << (3 * lvl)) - np.uint64(1)))
    keys = keys.astype(kd)
    ref = np.argsort(keys, kind="stable").astype(np.uint32)
    sk, order = gpu.sort_by_key(keys)
    assert np.array_equal(order, ref) and np.array_equal(sk, keys[ref])
    vals = rng.integers(0, 2 ** 31, n).astype(np.uint32)
    sk2, v2 = gpu.sort_by_key(keys, vals)
    assert np.array_equal(sk2, keys[ref]) and np.array_equal(v2, vals[ref])


def test_scan_and_minmax(gpu):
    import torch
    from cornerstone_octree_b200 import ops
    rng = np.random.default_rng(1)
    for n in (1, 7, 2048, 2049, 1000003):
        v = rng.integers(0, 1000, n).astype(np.uint32)
        d = torch.from_numpy(v.view(np.int32)).cuda()
        ex = ops.exclusive_scan(d, 5).cpu().numpy().view(np.uint32)
        inc = ops.inclusive_scan(d).cpu().numpy().view(np.uint32)
        cs = np.cumsum(v, dtype=np.uint64).astype(np.uint32)
        assert np.array_equal(inc, cs) and np.array_equal(ex, (cs - v + 5).astype(np.uint32))
    for fd in (np.float32, np.float64):
        x, y, z = gen.gaussian(100001, (-3, 2), fd)
        mm = ops.minmax3(*(torch.from_numpy(a).cuda() for a in (x, y, z)))
        assert list(mm) == [x.min(), x.max(), y.min(), y.max(), z.min(), z.max()]


def test_gather_multi_and_scatter(gpu):
    import torch
    from cornerstone_octree_b200 import ops
    rng = np.random.default_rng(2)
    n = 100001
    perm = rng.permutation(n).astype(np.int32)
    arrs = [rng.random(n), rng.random(n).astype(np.float32), rng.integers(0, 2 ** 62, n), rng.random(n)]
    d = [torch.from_numpy(a).cuda() for a in arrs]
    out = ops.gather_multi(torch.from_numpy(perm).cuda(), d)
    for a, o in zip(arrs, out):
        assert np.array_equal(o.cpu().numpy(), a[perm])
    back = torch.empty_like(d[0])
    ops.scatter(torch.from_numpy(perm).cuda(), out[0], back)
    assert np.array_equal(back.cpu().numpy(), arrs[0])


@pytest.mark.parametrize("narr", [1, 3, 4, 8, 9])
@pytest.mark.parametrize("n", [4095, 4096, 100003])
def test_gather_multi_all_double(narr, n, gpu):
    """the four-elements-per-thread kernel (8-byte arrays only, n >= 4096, tail of n % 4) and its fallbacks"""
    import torch
    from cornerstone_octree_b200 import ops
    rng = np.random.default_rng(narr * 7 + n)
    perm = rng.permutation(n).astype(np.int32)
    arrs = [rng.integers(0, 2 ** 63, n).view(np.float64) if k % 2 else rng.integers(0, 2 ** 63, n) for k in range(narr)]
    out = ops.gather_multi(torch.from_numpy(perm).cuda(), [torch.from_numpy(a).cuda() for a in arrs])
    for a, o in zip(arrs, out):
        assert np.array_equal(o.cpu().numpy().view(np.int64), a.view(np.int64)[perm])


@pytest.mark.parametrize("offset", [0.0, 1.0e6])
@pytest.mark.parametrize("bnd", [(0, 0, 0), (1, 1, 1)])
def test_neighbors_on_the_radius(offset, bnd, gpu, port):
    """Lattice particles with 2h an exact multiple of the spacing: many pairs sit EXACTLY on the search radius
    (excluded by the strict <) and many more within a few ulps of it.  The float32 filter of the double kernel
    must hand every such pair to the exact path; with `offset` the coordinates are large, so the filter's
    relative coordinates matter."""
    m = 24
    g = (np.arange(m) + 0.5) / m
    X, Y, Z = np.meshgrid(g, g, g, indexing="ij")
    rng = np.random.default_rng(3)
    perm = rng.permutation(m ** 3)
    x, y, z = (np.ascontiguousarray(a.ravel()[perm]) + offset for a in (X, Y, Z))
    h = np.full(x.size, 1.0 / m)           # 2h = 2 spacings: (2,0,0)-type pairs have d2 == r2 in exact arithmetic
    h[::3] = 0.5 * np.sqrt(5.0) / m        # r2 = 5 a^2: (2,1,0)-type pairs on the radius, inexact in binary
    box = loader.make_box((offset, offset + 1.0), bnd, np.float64)
    a = pipeline.run(port, 1, x, y, z, h, box, np.uint64, 16, 64)
    b = pipeline.run(gpu, 1, x, y, z, h, box, np.uint64, 16, 64)
    assert pipeline.compare(a, b, 64) == []
    assert a["neighbors_count"].min() >= 6


def test_neighbors_tiny_and_huge_h(gpu, port):
    """h far below float32 range relative to the box (exact-only lanes) and h of the order of the box"""
    x, y, z, h, box = make_case("gaussian", 8000, np.float64, (0, 0, 0))
    h[::2] = 1e-20
    h[1::7] = 0.4
    a = pipeline.run(port, 1, x, y, z, h, box, np.uint64, 16, 200)
    b = pipeline.run(gpu, 1, x, y, z, h, box, np.uint64, 16, 200)
    assert pipeline.compare(a, b, 200) == []


@pytest.mark.parametrize("bucket", [1, 2, 4])
@pytest.mark.parametrize("genname,bnd", [("gaussian", (0, 0, 0)), ("uniform", (1, 1, 1))])
def test_neighbors_merged_leaf_runs(bucket, genname, bnd, gpu, port):
    """tiny buckets: up to 8 sibling leaves (some empty) share one 32-candidate chunk of the neighbour search; every
    lane must keep exactly the candidates of the leaves its own walk reaches [neighbors.cu advanceChunk]"""
    x, y, z, h, box = make_case(genname, 6000, np.float64, bnd)
    a = pipeline.run(port, 1, x, y, z, h, box, np.uint64, bucket, 150)
    b = pipeline.run(gpu, 1, x, y, z, h, box, np.uint64, bucket, 150)
    assert pipeline.compare(a, b, 150) == []


def test_neighbors_count_only_and_plain_entry(gpu, port):
    """ngmax = 0 (no list, counts only, NULL neighbour pointer) and the entry point without workspace give the same
    counts / lists as the workspace path"""
    import torch
    from cornerstone_octree_b200 import ops
    import gpu_adapter as ga
    x, y, z, h, box = make_case("gaussian", 30000, np.float64, (0, 1, 0))
    a = pipeline.run(port, 1, x, y, z, h, box, np.uint64, 32, 64)
    oct_ = {k: a[k] for k in ("prefixes", "parents", "level_range", "internal_to_leaf", "leaf_to_internal")}
    oct_["child_offsets"] = np.concatenate((a["child_offsets"], np.zeros(1, np.int32)))
    L = a["tree"].size - 1
    oct_.update(num_leaf_nodes=L, num_internal_nodes=(L - 1) // 7, num_nodes=L + (L - 1) // 7)
    layout = np.zeros(L + 1, dtype=np.uint32)
    np.cumsum(a["counts"], out=layout[1:])
    d = ga.oct_to_dev(oct_)
    args = [ga.to_dev(a[k]) for k in ("x", "y", "z", "h")]
    n = x.size
    common = (0, n, ga.conv_box(box), d, ga.to_dev(layout), ga.to_dev(a["centers"]), ga.to_dev(a["sizes"]))
    # counts only
    nc0 = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    dummy = torch.empty((n, 0), dtype=torch.int32, device="cuda")
    ops.find_neighbors(*args, *common, 0, neighbors=dummy, neighbors_count=nc0)
    np.testing.assert_array_equal(nc0.cpu().numpy().view(np.uint32), a["neighbors_count"])
    # plain entry point (no workspace)
    nb1, nc1 = ops.find_neighbors(*args, *common, 64, use_workspace=False)
    nb1, nc1 = nb1.cpu().numpy().view(np.uint32), nc1.cpu().numpy().view(np.uint32)
    np.testing.assert_array_equal(nc1, a["neighbors_count"])
    for i in range(0, n, 97):
        k = min(int(nc1[i]), 64)
        np.testing.assert_array_equal(nb1[i, :k], a["neighbors"][i, :k])


@pytest.mark.parametrize("kd", [np.uint32, np.uint64])
def test_node_counts_unaligned_keys(kd, gpu, port):
    """particle keys that do not start on a 16-byte boundary take the plain-load kernel; same counts"""
    import torch
    from cornerstone_octree_b200 import ops
    import gpu_adapter as ga
    keys = np.sort(gen.gaussian_keys(100001, kd))
    tree, counts = port.compute_octree(keys[1:], 16)
    dk = ga.to_dev(keys)
    view = dk[1:]  # 4 or 8 bytes past a 256-byte aligned allocation
    assert view.data_ptr() % 16 != 0
    got = ops.compute_node_counts(ga.to_dev(tree), view)
    np.testing.assert_array_equal(got.cpu().numpy().view(np.uint32), counts)
    got = ops.compute_node_counts(ga.to_dev(tree), dk[2:] if kd == np.uint64 else dk[4:])
    want = port.compute_node_counts(tree, keys[2:] if kd == np.uint64 else keys[4:])
    np.testing.assert_array_equal(got.cpu().numpy().view(np.uint32), want)
