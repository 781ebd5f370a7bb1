"""the header-only device API include/cstone_b200/find_neighbors.cuh (cstone::traverseNeighbors with the
reference's signature, traversal/find_neighbors.cuh:435-446) inside a client kernel of the shape of the reference's
neighbour driver (tests/cpp/traverse_neighbors_client.cu).  Per particle: the count equals findNeighbors', the neighbour
SET equals findNeighbors' (order is the traversal's own, as in the reference), and what the client accumulated in place
from the per-warp scratch equals the same sum over findNeighbors' list."""
import ctypes as C
import math
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tests", "cpp", "_build", "libtraverse_client.so")


@pytest.mark.parametrize("bits,fd,pbc,n,bucket", [(64, "float64", False, 200000, 64), (64, "float64", True, 60000, 16),
                                                   (32, "float32", False, 100000, 64), (32, "float32", True, 30000, 8)])
def test_client_kernel_consumes_neighbours_in_place(bits, fd, pbc, n, bucket):
    if not os.path.exists(LIB):
        pytest.skip("tests/cpp/_build/libtraverse_client.so not built")
    import torch
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200.pipeline import HotPath
    lib = C.CDLL(LIB)
    tdt = getattr(torch, fd)
    ngmax = 150
    g = torch.Generator(device="cuda")
    g.manual_seed(11)
    x, y, z = (torch.rand(n, generator=g, device="cuda", dtype=torch.float64).to(tdt) for _ in range(3))
    hval = 0.5 * (3.0 * 60 / (4.0 * math.pi * n)) ** (1.0 / 3.0)
    h = (hval * (0.7 + 0.6 * torch.rand(n, generator=g, device="cuda", dtype=torch.float64))).to(tdt)
    bnd = (1, 1, 1) if pbc else (0, 0, 0)
    box = ops.make_box((0.0, 1.0), bnd, dtype=tdt)
    hp = HotPath(n, key_bits=bits, fdtype=tdt, bucket=bucket, ngmax=ngmax)
    L = hp.step(x, y, z, h, box, cold=True)
    torch.cuda.synchronize()
    xs, ys, zs, hs = hp.sorted
    v = hp.octree
    first, last = n // 7, n - n // 5  # a window of targets, like Domain's [startIndex, endIndex)
    nw = last - first
    blocks = 148 * 4
    nidx_n, pool_n = C.c_uint64(), C.c_uint64()
    lib.traverse_client_pool_sizes(blocks, ngmax, C.byref(nidx_n), C.byref(pool_n))
    nidx = torch.empty(nidx_n.value, dtype=torch.int32, device="cuda")
    pool = torch.empty(pool_n.value, dtype=torch.int32, device="cuda")
    nc = torch.zeros(nw, dtype=torch.int32, device="cuda")
    nb = torch.zeros((nw, ngmax), dtype=torch.int32, device="cuda")
    hsum = torch.zeros(nw, dtype=torch.float64, device="cuda")
    fn = getattr(lib, "traverse_client_k%d_f%s" % (bits, fd[-2:]))
    p = lambda t: C.c_void_p(t.data_ptr())
    st = fn(p(xs), p(ys), p(zs), p(hs), first, last, L, p(v.prefixes), p(v.child_offsets), p(v.internal_to_leaf),
            p(v.level_range), p(hp.leaves.tree_a), p(hp.layout), p(hp.centers), p(hp.sizes), C.byref(box), ngmax, p(nidx),
            p(pool), blocks, p(nc), p(nb), p(hsum))
    assert st == 0
    # the checker: findNeighbors of the library over the same window (bit-exact against the reference elsewhere)
    rnb, rnc = ops.find_neighbors(xs, ys, zs, hs, first, last, box, v, hp.layout[:L + 1], hp.centers, hp.sizes, ngmax)
    nc_h, rnc_h = nc.cpu().numpy(), rnc.cpu().numpy()
    assert np.array_equal(nc_h, rnc_h)
    assert 20 < nc_h.mean() < ngmax
    keep = np.arange(ngmax)[None, :] < np.minimum(nc_h, ngmax)[:, None]
    a = np.where(keep, nb.cpu().numpy().astype(np.int64), -1)
    b = np.where(keep, rnb.cpu().numpy().astype(np.int64), -1)
    full = nc_h <= ngmax
    assert np.array_equal(np.sort(a[full], axis=1), np.sort(b[full], axis=1))
    # consumed in place: the client's sum over the warp scratch against the sum over the particle-major list
    hs_h = hs.cpu().numpy().astype(np.float64)
    ref_sum = np.where(keep, hs_h[np.where(keep, a, 0)], 0.0).sum(axis=1)
    assert np.allclose(hsum.cpu().numpy()[full], ref_sum[full], rtol=1e-12, atol=0)
