"""GPU probe: per-stage times of the single-rank hot path (cornerstone_octree_b200.pipeline.HotPath) for float32 and
float64 coordinates on the same uniform cloud.  usage: nb_probe.py [n] [bucket]"""
import math, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cornerstone_octree_b200 import ops
from cornerstone_octree_b200.pipeline import HotPath

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 64
g = torch.Generator(device="cuda"); g.manual_seed(3)
x64, y64, z64 = (torch.rand(n, generator=g, device="cuda", dtype=torch.float64) for _ in range(3))
hval = 0.5 * (3.0 * 100 / (4.0 * math.pi * n)) ** (1.0 / 3.0)
for fd, bits in ((torch.float64, 64), (torch.float32, 64), (torch.float32, 32)):
    x, y, z = x64.to(fd), y64.to(fd), z64.to(fd)
    h = torch.full((n,), hval, device="cuda", dtype=fd)
    box = ops.make_box((0.0, 1.0), dtype=fd)
    hp = HotPath(n, key_bits=bits, fdtype=fd, bucket=bucket, ngmax=150)
    for it in range(3):
        L = hp.step(x, y, z, h, box, cold=(it == 0), time_stages=True)
    torch.cuda.synchronize()
    ms = hp.stage_ms()
    print("%s k%d n=%d leaves=%d nc=%.1f | " % (str(fd).split(".")[1], bits, n, L, hp.neighbors_count.float().mean().item()) +
          " ".join("%s %.3f" % (k, v) for k, v in ms.items()))
    del hp
