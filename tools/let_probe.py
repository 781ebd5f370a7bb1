"""Diagnostic: halo mode 1 (multi-rank focus tree) against the reference Domain, stage by stage, without asserting.
usage: python tools/let_probe.py [cloud] [P] [n] [bucket] [bucket_focus] [syncs]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import test_gpu_domain_ranks as T  # noqa: E402


def first_diff(a, b):
    n = min(a.size, b.size)
    d = np.nonzero(a[:n] != b[:n])[0]
    return "sizes %d/%d first diff at %s" % (a.size, b.size, d[0] if d.size else "-")


def main():
    a = sys.argv[1:]
    cloud = a[0] if len(a) > 0 else "gaussian"
    P, n, bucket, bf = (int(a[i]) if len(a) > i else d for i, d in ((1, 2), (2, 20000), (3, 64), (4, 8)))
    syncs = int(a[5]) if len(a) > 5 else 3
    x, y, z, h, lim, pbc = T.make_cloud(cloud, n)
    rk = T.Ranks(P, bucket, bf, lim, pbc, x, y, z, h, halo_mode=1)
    for it in range(syncs):
        if it:
            rk.displace(0.01)
        try:
            rk.sync()
        except Exception as e:  # noqa: BLE001
            print("sync %d raised: %s" % (it, e))
            break
        for r in range(P):
            dom, inf, tr = rk.doms[r], rk.ref.info(r), rk.ref.trees(r)
            lv = dom.let_view()
            ok = True
            for name, mine, ref in (("leaves", lv["leaves"], tr["focus_leaves"]), ("counts", lv["counts"], tr["focus_counts"]),
                                    ("assignment", lv["assignment"].ravel(), tr["focus_assignment"].ravel()),
                                    ("layout", lv["layout"], tr["layout"])):
                same = np.array_equal(mine, ref)
                ok = ok and same
                if not same:
                    print("sync %d rank %d %s DIFFER: %s" % (it, r, name, first_diff(np.asarray(mine), np.asarray(ref))))
            rng = (dom.start_index, dom.end_index, dom.n_particles_with_halos)
            rref = (inf["start"], inf["end"], inf["size"])
            print("sync %d rank %d: L %d peers %s ranges %s ref %s flags %d %s" % (
                it, r, lv["leaves"].size - 1, lv["peers"].tolist(), rng, rref, int(lv["flags"].sum()),
                "OK" if ok and rng == rref else "MISMATCH"))
        try:
            rk.check()
            print("sync %d: full check passed" % it)
        except AssertionError as e:
            print("sync %d: check failed: %s" % (it, str(e)[:300]))
    rk.close()


if __name__ == "__main__":
    main()
