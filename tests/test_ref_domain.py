"""Pin the multi-rank oracle: the reference's Domain<KeyType,T,CpuTag> on P thread-ranks (oracle/ref_domain.cpp over
oracle/mpi_shim) must reproduce the invariants the reference's own integration test asserts
[ref: test/integration_mpi/domain_nranks.cpp:113-149]: particle conservation, keys equal to computeSfcKeys of the
returned coordinates under the returned box, keys sorted, and the sum of neighbour counts over all ranks equal to an
all-to-all search over the global cloud [ref: test/unit/neighbors/all_to_all.hpp]."""
import numpy as np
import pytest

from oracle import loader
import gen

pytestmark = pytest.mark.skipif(not loader.refdom_available(),
                                reason="oracle/_ref/liboracle_refdom.so not built (reference tree absent)")


def gaussian_cloud(n, fd=np.float64):
    """domain_nranks.cpp:57-101: Gaussian in [-1,1]^3, h grows with the distance from the centre"""
    x, y, z = gen.gaussian(n, (-1.0, 1.0), fd)
    h = (0.05 + 0.2 * (x * x + y * y + z * z)).astype(fd)
    return x, y, z, h


def all_to_all_counts(x, y, z, h, pbc=None):
    """[ref: test/unit/neighbors/all_to_all.hpp] count of j != i with |r_j - r_i|^2 < 4 h_i^2"""
    n = x.size
    out = np.zeros(n, dtype=np.int64)
    for s in range(0, n, 512):
        e = min(n, s + 512)
        d2 = np.zeros((e - s, n))
        for d, c in enumerate((x, y, z)):
            dd = c[None, :].astype(np.float64) - c[s:e, None].astype(np.float64)
            if pbc is not None and pbc[d]:
                dd -= pbc[d] * np.rint(dd / pbc[d])
            d2 += dd * dd
        r2 = (2.0 * h[s:e].astype(np.float64)) ** 2
        out[s:e] = (d2 < r2[:, None]).sum(axis=1) - 1
    return out


@pytest.mark.parametrize("P", [1, 2, 4, 8])
def test_reference_invariants(P, ref):
    n = (4000 // P) * P
    x, y, z, h = gaussian_cloud(n)
    rd = loader.RefDomain(P, 50, 10, 0.5, [-1, 1])
    for r in range(P):
        s = slice(r * n // P, (r + 1) * n // P)
        rd.set(r, x[s], y[s], z[s], h[s])
    for step in range(2):  # the second sync starts from the first one's trees and exchanges almost nothing
        rd.sync()
        total, nsum = 0, 0
        nbs = rd.find_neighbors(300)
        for r in range(P):
            inf = rd.info(r)
            keys, rx, ry, rz, rh = rd.arrays(r)
            assert keys.size == inf["size"]
            total += inf["end"] - inf["start"]
            box = loader.make_box(inf["box"])
            assert np.array_equal(keys, ref.compute_sfc_keys(1, rx, ry, rz, box, np.uint64))
            assert np.all(keys[1:] >= keys[:-1])
            tr = rd.trees(r)
            assert tr["layout"][inf["start_cell"]] == inf["start"] and tr["layout"][inf["end_cell"]] == inf["end"]
            assert tr["layout"][-1] == inf["size"]
            nsum += int(nbs[r][1].sum())
        assert total == n
        assert nsum == int(all_to_all_counts(x, y, z, h).sum())
    rd.close()


def test_properties_and_halo_exchange(ref):
    """conserved properties follow their particles; exchangeHalos fills the halo ranges with the owner's values"""
    P, n = 4, 8000
    x, y, z = gen.uniform(n, (0.0, 1.0))
    h = np.full(n, 0.5 * np.cbrt(3 * 40 / (4 * np.pi * n)))
    ident = np.arange(n, dtype=np.float64)
    rd = loader.RefDomain(P, 64, 8, 0.5, [0, 1], pbc=(1, 1, 1), nprops=1)
    for r in range(P):
        s = slice(r * n // P, (r + 1) * n // P)
        rd.set(r, x[s], y[s], z[s], h[s], props=[ident[s]])
    rd.sync()
    seen, halo_in = [], []
    for r in range(P):
        inf = rd.info(r)
        keys, rx, ry, rz, rh, rid = rd.arrays(r)
        a, e = inf["start"], inf["end"]
        ids = rid[a:e].astype(np.int64)
        assert np.array_equal(rx[a:e], x[ids]) and np.array_equal(ry[a:e], y[ids]) and np.array_equal(rz[a:e], z[ids])
        seen.append(ids)
        buf = np.full(inf["size"], -1.0)
        buf[a:e] = rid[a:e]
        halo_in.append(buf)
    assert np.array_equal(np.sort(np.concatenate(seen)), np.arange(n))
    rd.exchange_halos(halo_in)
    nsum = 0
    for r in range(P):
        inf = rd.info(r)
        keys, rx, ry, rz, rh, rid = rd.arrays(r)
        ids = halo_in[r].astype(np.int64)
        halo = np.ones(inf["size"], dtype=bool)
        halo[inf["start"]:inf["end"]] = False
        filled = halo & (ids >= 0)  # halo slots of leaves without particles on the owner stay untouched
        assert np.array_equal(rx[filled], x[ids[filled]]) and np.array_equal(rz[filled], z[ids[filled]])
    for nb, nc in rd.find_neighbors(200):
        nsum += int(nc.sum())
    assert nsum == int(all_to_all_counts(x, y, z, h, pbc=(1.0, 1.0, 1.0)).sum())
    rd.close()
