"""GPU probe: P engine instances as P host threads on ONE GPU (in-process transport), steady-state syncs with drift —
the multi-rank kernels (merge re-sort, two-source gather, surface boxes, reverse collision, send lists, halo pack) under
one process, e.g. for ncu.  usage: local_ranks_probe.py [P] [n_per_rank] [steps] [halo_mode]"""
import math, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from cornerstone_octree_b200 import ops
from cornerstone_octree_b200.domain import Domain, LocalWorld
P = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8_000_000
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
halo_mode = int(sys.argv[4]) if len(sys.argv) > 4 else 0
world = LocalWorld(P)
doms = [Domain(r, P, 65536, 64, 0.5, ops.make_box((0.0, 1.0))) for r in range(P)]
for d in doms:
    d.set_local_world(world)
    if halo_mode:
        d.set_halo_mode(halo_mode)
state = []
for r in range(P):
    x, y, z, h, _ = bench.make_particles(n, "uniform", 42 + r)
    h = h / P ** (1.0 / 3.0)
    state.append([torch.from_numpy(a).cuda() for a in (x, y, z, h)])
amp = 0.1 * float(state[0][3][0])
for s in range(steps):
    sign = 1.0 if s % 2 == 0 else -1.0
    def one(r):
        xx, yy, zz, hh = state[r]
        w = 2.0 * math.pi * 37.0
        xx.add_(torch.sin(yy * w), alpha=amp * sign); yy.add_(torch.sin(zz * w), alpha=amp * sign); zz.add_(torch.sin(xx * w), alpha=amp * sign)
        keys, sx, sy, sz, sh, _ = doms[r].sync(xx, yy, zz, hh)
        state[r] = [sx, sy, sz, sh]
        return doms[r].stage_ms()
    ms = world.run(one)
    torch.cuda.synchronize()
    print("step", s, {k: round(v, 3) for k, v in ms[0].items()}, "halos", doms[0].n_particles_with_halos - doms[0].n_particles, flush=True)
