import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import gen, pipeline, gpu_adapter
from oracle import loader
port = loader.load("port")
gpu = gpu_adapter.GpuImpl()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1200
nbad = 0
for kd in (np.uint32, np.uint64):
    for limits in ((0., 1., 0., 1., 0., 1.), (-1.2, 0.23, -0.213, 3.213, -5.1, 1.23)):
        for pbc in (0, 1):
            for cloud in ("uniform", "gaussian"):
                for radius in (0.124, 0.0624):
                    box = loader.make_box(limits, (pbc,) * 3, np.float64)
                    x, y, z = (gen.uniform if cloud == "uniform" else gen.gaussian)(n, limits, np.float64)
                    h = np.full(n, radius / 2)
                    a = pipeline.run(port, 1, x, y, z, h, box, kd, 64, n, halo_range=(0, 1))
                    b = pipeline.run(gpu, 1, x, y, z, h, box, kd, 64, n, halo_range=(0, 1))
                    bad = pipeline.compare(a, b, n)
                    nc_a, nc_b = a["neighbors_count"], b["neighbors_count"]
                    diff = np.nonzero(nc_a != nc_b)[0]
                    print(kd.__name__, limits[1], pbc, cloud, radius, "bad:", bad, "leaves", a["tree"].size - 1,
                          "count diffs", diff.size, diff[:8], flush=True)
                    for i in diff[:3]:
                        sa = set(a["neighbors"][i, :nc_a[i]].tolist()); sb = set(b["neighbors"][i, :nc_b[i]].tolist())
                        print("   i", i, "oracle-only", sorted(sa - sb)[:10], "gpu-only", sorted(sb - sa)[:10])
                        lay = np.concatenate([[0], np.cumsum(a["counts"])])
                        for j in sorted(sa ^ sb)[:3]:
                            lj = np.searchsorted(lay, j, side="right") - 1
                            li = np.searchsorted(lay, i, side="right") - 1
                            print("      j", j, "leaf(j)", lj, "range", lay[lj], lay[lj + 1], "leaf(i)", li, lay[li], lay[li + 1])
                    nbad += len(bad) > 0
print("configs with mismatch:", nbad)
