"""GPU probe: sort_by_key on a small, digit-skewed key set (shape of the octree node sort in buildOctree)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cornerstone_octree_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_400_000
g = torch.Generator(device="cuda"); g.manual_seed(3)
lo = torch.randint(0, 1 << 24, (n,), device="cuda", dtype=torch.int64, generator=g)
cases = {"low24": lo,
         "low24+top2": lo | (torch.randint(0, 4, (n,), device="cuda", dtype=torch.int64, generator=g) << 60),
         "full63": torch.randint(0, 1 << 62, (n,), device="cuda", dtype=torch.int64, generator=g)}
for name, k0 in cases.items():
    srt = ops.SfcSorter("cuda")
    srt.order = torch.arange(n, device="cuda", dtype=torch.int32)[1:]  # misaligned-length payload path is not needed
    srt.order = None
    keys = k0.clone()
    order = srt.set_map_from_codes(keys)
    rk, ro = torch.sort(k0, stable=True)
    ok = bool(torch.equal(keys, rk)) and bool(torch.equal(order.long(), ro))
    ms = []
    for _ in range(10):
        keys.copy_(k0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); srt.update_map(keys); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    print("%-12s n=%d parity=%s  %.3f ms (min %.3f)" % (name, n, ok, ms[5], ms[0]))
