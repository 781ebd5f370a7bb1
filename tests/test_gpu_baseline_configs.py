"""Bit-exact GPU parity at the BASELINE.json configurations themselves:
 c1  1 M uniform particles, u64 Hilbert, bucket 64, ngmax 150 — every stage incl. the raw neighbour lists, against the
     reference's own CPU code (oracle/_ref; the port if it did not travel);
 c5  u32 MORTON keys, bucket 16, float and double: keys, sort, tree, linking, centres, findHalos, findNeighbors on a
     Morton tree, incl. leaves that saturate at level 10 and keep more than `bucket` particles
     [ref: tree/csarray.hpp:303-306].  The reference hard-wires Hilbert in nodeFpCenters and findHalos
     (sfc/sfc.hpp:54-55, focus/source_center.hpp:154, traversal/collisions.hpp:51), so the checker of those two
     stages on a Morton tree is the port's restatement with mortonIBox [ref: sfc/morton.hpp:175-182]; keys, sort, leaf
     array and linking are ALSO compared with the reference build;
 a2  keys equal to removeKey are preserved by key generation [ref: sfc/sfc.hpp:283-291, tree/definitions.h:86-91]."""
import numpy as np
import pytest

import gen
import pipeline
from oracle import loader

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import gpu_adapter
    return gpu_adapter.GpuImpl()


def test_c1_one_million_uniform_bit_exact(gpu, oracle):
    """BASELINE configs[0]: computeOctree + findNeighbors, 1 M uniform, 64-bit Hilbert, bucketSize 64"""
    n = 1_000_000
    x, y, z = gen.uniform(n, (0.0, 1.0), np.float64)
    h = np.full(n, 0.5 * np.cbrt(3 * 100 / (4 * np.pi * n)))
    box = loader.make_box((0.0, 1.0), (0, 0, 0), np.float64)
    a = pipeline.run(oracle, 1, x, y, z, h, box, np.uint64, 64, 150)
    b = pipeline.run(gpu, 1, x, y, z, h, box, np.uint64, 64, 150)
    assert a["tree"].size - 1 == 32768  # verified with the reference
    assert 90 < a["neighbors_count"].mean() < 110
    assert pipeline.compare(a, b, 150) == []


@pytest.mark.parametrize("fd", [np.float32, np.float64])
@pytest.mark.parametrize("genname,n,bnd", [("uniform", 60000, (0, 0, 0)), ("gaussian", 60000, (1, 0, 1))])
def test_c5_morton_u32_bucket16(fd, genname, n, bnd, gpu, port):
    """BASELINE configs[4] arithmetic: 32-bit Morton keys, bucketSize 16 — whole pipeline on the Morton tree"""
    lim = (0.0, 1.0) if genname == "uniform" else (-1.0, 1.0)
    x, y, z = getattr(gen, genname)(n, lim, fd)
    box = loader.make_box(lim, bnd, fd)
    h = np.full(n, 0.5 * (lim[1] - lim[0]) * np.cbrt(3 * 40 / (4 * np.pi * n)), dtype=fd)
    a = pipeline.run(port, loader.MORTON, x, y, z, h, box, np.uint32, 16, 100)
    b = pipeline.run(gpu, loader.MORTON, x, y, z, h, box, np.uint32, 16, 100)
    assert pipeline.compare(a, b, 100) == []
    assert a["flags"].sum() > 0 and a["neighbors_count"].mean() > 10
    if loader.available("ref"):  # the stages the reference itself supports on Morton keys
        r = loader.load("ref")
        keys = r.compute_sfc_keys(loader.MORTON, x, y, z, box, np.uint32)
        sk, order = r.sort_by_key(keys)
        tree, counts = r.compute_octree(sk, 16)
        oc = r.build_octree(tree)
        assert np.array_equal(keys, b["keys_unsorted"]) and np.array_equal(order, b["order"])
        assert np.array_equal(tree, b["tree"]) and np.array_equal(counts, b["counts"])
        for k in ("prefixes", "parents", "level_range", "internal_to_leaf", "leaf_to_internal"):
            assert np.array_equal(oc[k], b[k]), k


def test_c5_morton_u32_saturates_at_level_10(gpu, port):
    """a tight cluster inside one level-10 cell: its leaf cannot split further and keeps count > bucket"""
    n = 30000
    fd = np.float64
    x, y, z = gen.uniform(n, (0.0, 1.0), fd)
    cell = 1.0 / 1024
    rng = np.random.default_rng(5)
    m = 400  # 400 particles in ONE level-10 cell, bucket 16
    for c in (x, y, z):
        c[:m] = (517 + 0.05 + 0.9 * rng.random(m)) * cell
    box = loader.make_box((0.0, 1.0), (0, 0, 0), fd)
    h = np.full(n, 0.5 * np.cbrt(3 * 30 / (4 * np.pi * n)))
    h[:m] = 0.4 * cell
    a = pipeline.run(port, loader.MORTON, x, y, z, h, box, np.uint32, 16, 500)
    b = pipeline.run(gpu, loader.MORTON, x, y, z, h, box, np.uint32, 16, 500)
    assert a["counts"].max() >= m
    assert pipeline.compare(a, b, 500) == []


@pytest.mark.parametrize("kd,fd,kind", [(np.uint64, np.float64, 1), (np.uint32, np.float32, 0), (np.uint64, np.float32, 0),
                                        (np.uint32, np.float64, 1)])
def test_remove_key_is_preserved_on_gpu(kd, fd, kind, gpu, oracle):
    """keys[i] == removeKey on input stay removeKey on output, every other key is computed; removed particles then
    sort to the end"""
    n = 100003
    x, y, z = gen.uniform(n, (0.0, 1.0), fd)
    box = loader.make_box((0.0, 1.0), (0, 0, 0), fd)
    remove = np.array(1 << (3 * loader.max_level(kd)), dtype=kd)
    rng = np.random.default_rng(11)
    keys_in = rng.integers(0, 1 << 20, n).astype(kd)  # arbitrary stale content
    marked = rng.random(n) < 0.03
    keys_in[marked] = remove
    a = oracle.compute_sfc_keys(kind, x, y, z, box, kd, keys=keys_in.copy())
    b = gpu.compute_sfc_keys(kind, x, y, z, box, kd, keys=keys_in.copy())
    assert np.array_equal(a, b)
    assert np.all(b[marked] == remove) and not np.any(b[~marked] == remove)
    sk, order = gpu.sort_by_key(b)
    assert np.all(sk[-int(marked.sum()):] == remove) and np.all(marked[order[-int(marked.sum()):]])
