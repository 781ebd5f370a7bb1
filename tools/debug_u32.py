"""debug: steady-state Domain (u32 keys, bucket 16) neighbour counts against all-pairs; prints the first discrepancy"""
import math, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from cornerstone_octree_b200 import ops
from cornerstone_octree_b200.domain import Domain
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
bits = int(sys.argv[2]) if len(sys.argv) > 2 else 32
bucket = int(sys.argv[3]) if len(sys.argv) > 3 else 16
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 4
x, y, z, h, lim = bench.make_particles(n, "uniform", 42)
dom = Domain(0, 1, 65536, bucket, 0.5, ops.make_box(lim), key_bits=bits)
a = [torch.from_numpy(v).cuda() for v in (x, y, z, h)]
amp = 0.1 * h[0]; sign = 1.0
for s in range(steps):
    w = 2.0 * math.pi * 37.0
    a[0].add_(torch.sin(a[1] * w), alpha=amp * sign); a[1].add_(torch.sin(a[2] * w), alpha=amp * sign); a[2].add_(torch.sin(a[0] * w), alpha=amp * sign)
    sign = -sign
    keys, sx, sy, sz, sh, _ = dom.sync(*a)
    a = [sx, sy, sz, sh]
    nb, nc = dom.find_neighbors(sx, sy, sz, sh, 150)
    idx = torch.arange(256, device="cuda") * (n - 1) // 255
    bad = 0
    for q in idx.tolist():
        d2 = (sx - sx[q]) ** 2 + (sy - sy[q]) ** 2 + (sz - sz[q]) ** 2
        exp = int((d2 < 4.0 * sh[q] * sh[q]).sum().item()) - 1
        got = int(nc[q].item())
        if exp != got:
            bad += 1
            if bad == 1:
                hit = torch.nonzero(d2 < 4.0 * sh[q] * sh[q]).flatten()
                lst = set(nb[q, :min(got, 150)].tolist())
                miss = [j for j in hit.tolist() if j != q and j not in lst]
                box = dom.box()
                print("step", s, "target", q, "expected", exp, "got", got, "missing", miss[:5], "box", [box.limits[i] for i in range(6)])
                t = dom.tree_arrays()
                leaves = t["leaves"].cpu().numpy().view(np.uint32 if bits == 32 else np.uint64)
                lay = t["layout"].cpu().numpy()
                k = keys.cpu().numpy().view(leaves.dtype)
                for j in miss[:3]:
                    leaf = np.searchsorted(lay, j, side="right") - 1
                    print("   j", j, "key", hex(k[j]), "leaf", leaf, "range", hex(leaves[leaf]), hex(leaves[leaf + 1]), "layout", lay[leaf], lay[leaf + 1],
                          "xyz", sx[j].item(), sy[j].item(), sz[j].item(), "key in leaf:", leaves[leaf] <= k[j] < leaves[leaf + 1])
    if bad:
        # the reference's own findNeighbors and node centres on the SAME arrays and tree
        from oracle import loader
        ref = loader.load("ref")
        t = dom.tree_arrays()
        kd = np.uint32 if bits == 32 else np.uint64
        oc = {k: t[k].cpu().numpy() for k in ("child_offsets", "internal_to_leaf")}
        lay = t["layout"].cpu().numpy().view(np.uint32)
        b = dom.box()
        rbox = loader.make_box([b.limits[i] for i in range(6)])
        cen, siz = t["centers"].cpu().numpy().reshape(-1, 3), t["sizes"].cpu().numpy().reshape(-1, 3)
        rc, rs = ref.geo_centers(1, t["prefixes"].cpu().numpy().view(kd), rbox, np.float64)
        print("   centres equal reference:", np.array_equal(cen, rc), np.array_equal(siz, rs))
        hx, hy, hz, hh = (v.cpu().numpy() for v in (sx, sy, sz, sh))
        rnb, rnc = ref.find_neighbors(hx, hy, hz, hh, 0, n, rbox, oc, lay, cen, siz, 150)
        g = nc.cpu().numpy().view(np.uint32)
        print("   GPU counts equal reference findNeighbors on the same tree:", np.array_equal(g, rnc))
        rk = ref.compute_sfc_keys(1, hx, hy, hz, rbox, kd)
        print("   keys equal reference keys:", np.array_equal(keys.cpu().numpy().view(kd), rk))
    print("step", s, "mismatches", bad, "of 256; box", [dom.box().limits[i] for i in range(6)], flush=True)
