"""GPU probe: the neighbour search alone on the tree of one HotPath step (float64, u64 Hilbert).
usage: nb_only.py [n] [bucket] [reps] [plummer]   env CSB_NB_WALK=1: warp walk instead of the block-staged kernel"""
import math, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cornerstone_octree_b200 import ops
from cornerstone_octree_b200.pipeline import HotPath

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 64
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
g = torch.Generator(device="cuda"); g.manual_seed(3)
fd = torch.float64
x, y, z = (torch.rand(n, generator=g, device="cuda", dtype=fd) for _ in range(3))
hval = 0.5 * (3.0 * 100 / (4.0 * math.pi * n)) ** (1.0 / 3.0)
h = torch.full((n,), hval, device="cuda", dtype=fd)
box = ops.make_box((0.0, 1.0), dtype=fd)
hp = HotPath(n, key_bits=64, fdtype=fd, bucket=bucket, ngmax=150)
for it in range(2):
    L = hp.step(x, y, z, h, box, cold=(it == 0), time_stages=True)
torch.cuda.synchronize()
xs, ys, zs, hs = hp.sorted
ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
ev[0].record()
for r in range(reps):
    ops.find_neighbors(xs, ys, zs, hs, 0, n, box, hp.octree, hp.layout[:L + 1], hp.centers, hp.sizes, 150, hp.neighbors,
                       hp.neighbors_count)
    ev[r + 1].record()
torch.cuda.synchronize()
print("n=%d leaves=%d nc=%.2f ms/call: %s" % (n, L, hp.neighbors_count.float().mean().item(),
      " ".join("%.3f" % ev[r].elapsed_time(ev[r + 1]) for r in range(reps))))
