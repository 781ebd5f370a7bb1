"""GPU probe: gather_multi time for identity / locally shuffled / random permutations (4 double arrays, 32 M)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cornerstone_octree_b200 import ops
n = 32_000_000
src = [torch.rand(n, dtype=torch.float64, device="cuda") for _ in range(4)]
dst = [torch.empty_like(s) for s in src]
i = torch.arange(n, device="cuda", dtype=torch.float64)
perms = {"identity": torch.arange(n, device="cuda", dtype=torch.int32),
         "local~50": torch.argsort(i + 50.0 * torch.rand(n, device="cuda", dtype=torch.float64)).to(torch.int32),
         "local~500": torch.argsort(i + 500.0 * torch.rand(n, device="cuda", dtype=torch.float64)).to(torch.int32),
         "random": torch.randperm(n, device="cuda").to(torch.int32)}
for name, p in perms.items():
    for _ in range(2):
        ops.gather_multi(p, src, dst)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ops.gather_multi(p, src, dst)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("%-10s %.3f ms  %.0f GB/s algorithmic" % (name, ms, (n * (4 + 64)) / ms / 1e6))
