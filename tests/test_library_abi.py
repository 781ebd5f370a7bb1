"""CPU-side checks of the product library: it loads, exports every symbol include/cstone_b200.h declares, and its
host-generated Hilbert state-machine tables reproduce the oracle's curve (no compute calls, no GPU needed)."""
import ctypes as C
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_match_header():
    import cornerstone_octree_b200 as pkg
    header = open(os.path.join(ROOT, "include", "cstone_b200.h")).read()
    declared = set(re.findall(r"\b(cstone_b200_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 50
    lib = C.CDLL(pkg.LIB_PATH)
    missing = [name for name in sorted(declared) if not hasattr(lib, name)]
    assert missing == []
    assert lib.cstone_b200_abi_version() == 1
    # python prototypes cover every declared compute entry point
    bound = {"cstone_b200_" + n for n in pkg._lib.EXPORTED}
    assert declared - bound == set()


def test_status_strings():
    import cornerstone_octree_b200 as pkg
    assert pkg._lib.status_string(0) == "success"
    assert "capacity" in pkg._lib.status_string(-5)


def test_no_oracle_in_product():
    """the product package must not import or link the CPU checkers"""
    pkgdir = os.path.join(ROOT, "cornerstone_octree_b200")
    for dirpath, _, files in os.walk(pkgdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


def test_hilbert_tables_match_oracle(port):
    import cornerstone_octree_b200 as pkg
    lib = pkg._lib.lib
    enc = np.ctypeslib.as_array(lib.cstone_b200_host_hilbert_table(0), shape=(24 * 512,)).copy()
    dec = np.ctypeslib.as_array(lib.cstone_b200_host_hilbert_table(1), shape=(24 * 512,)).copy()

    def encode(ix, iy, iz, ml):
        steps = (ml + 2) // 3
        pad = steps * 3 - ml
        ix, iy, iz = ix << pad, iy << pad, iz << pad
        key = st = 0
        for s in range(steps - 1, -1, -1):
            idx = (((ix >> (3 * s)) & 7) << 6) | (((iy >> (3 * s)) & 7) << 3) | ((iz >> (3 * s)) & 7)
            e = int(enc[st * 512 + idx])
            key, st = (key << 9) | (e & 511), e >> 9
        return key >> (3 * pad)

    def decode(key, levels, ml):
        steps = (ml + 2) // 3
        pad = steps * 3 - ml
        k = key << (3 * pad)
        st = x = y = z = 0
        for s in range(steps - 1, -1, -1):
            if steps - 1 - s < (levels + 2) // 3:
                e = int(dec[st * 512 + ((k >> (9 * s)) & 511)])
                st = e >> 9
                x |= ((e >> 6) & 7) << (3 * s)
                y |= ((e >> 3) & 7) << (3 * s)
                z |= (e & 7) << (3 * s)
        mask = ~((1 << (ml - levels)) - 1)
        return ((x >> pad) & mask, (y >> pad) & mask, (z >> pad) & mask)

    rng = np.random.default_rng(0)
    for kd, ml in ((np.uint32, 10), (np.uint64, 21)):
        for _ in range(1500):
            p = [int(v) for v in rng.integers(0, 1 << ml, 3)]
            k = int(port.encode(1, *p, kd))
            assert encode(*p, ml) == k and decode(k, ml, ml) == tuple(p)
            lvl = int(rng.integers(0, ml + 1))
            start = (k >> (3 * (ml - lvl))) << (3 * (ml - lvl))
            b = port.sfc_ibox(1, start, lvl, kd)
            assert decode(start, lvl, ml) == (b[0], b[2], b[4])
