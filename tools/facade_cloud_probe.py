"""Diagnostic: the cloud of tests/cpp/domain_facade_test.cpp (tools/facade_cloud_6000.bin) on 2 thread-ranks, halo mode 1:
two syncs, then syncGrav twice (error tracing on)."""
import os, sys
os.environ["CSB_TRACE_ERRORS"] = "1"
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cornerstone_octree_b200 import ops
from cornerstone_octree_b200.domain import Domain, LocalWorld
n, P = 6000, 2
a = np.fromfile(os.path.join(ROOT, "tools", "facade_cloud_6000.bin"), dtype=np.float64).reshape(4, n)
world = LocalWorld(P)
doms = [Domain(r, P, 64, 8, 1.0, ops.make_box((-1.0, 1.0))) for r in range(P)]
state = []
for r, d in enumerate(doms):
    d.set_local_world(world); d.set_halo_mode(1)
    s = slice(r * n // P, (r + 1) * n // P)
    state.append([torch.from_numpy(a[k][s].copy()).cuda() for k in range(4)] + [torch.full((n // P,), 1.0 / n, dtype=torch.float64).cuda()])
for it in range(4):
    grav = it >= 2
    def one(r):
        sx, sy, sz, sh, sm = state[r]
        try:
            if grav:
                k, sx, sy, sz, sh, sm, _ = doms[r].sync_grav(sx, sy, sz, sh, sm)
            else:
                k, sx, sy, sz, sh, (sm,) = doms[r].sync(sx, sy, sz, sh, [sm])
        except Exception as e:
            print("rank", r, "step", it, "raised", e, flush=True)
            os._exit(3)
        state[r] = [sx, sy, sz, sh, sm]
    world.run(one)
    for r in range(P):
        lv = doms[r].let_view()
        print("step", it, "grav" if grav else "sync", "rank", r, "L", lv["leaves"].size - 1, "peers", lv["peers"].tolist(),
              "assignment", lv["assignment"].tolist(), (doms[r].start_index, doms[r].end_index, doms[r].n_particles_with_halos), flush=True)
print("done")
