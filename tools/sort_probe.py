"""GPU probe: sort_by_key (onesweep) time and correctness against torch.sort(stable) — 64-bit and 32-bit keys."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cornerstone_octree_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
g = torch.Generator(device="cuda"); g.manual_seed(7)
for name, dt, bits, passes in (("u64", torch.int64, 63, 8), ("u64-nearly-sorted", torch.int64, 63, 8),
                               ("u32", torch.int32, 30, 4)):
    if name == "u64-nearly-sorted":  # what a time-stepping code feeds the sort: sorted keys with perturbed low bits
        k0 = torch.sort(k0)[0] ^ torch.randint(0, 1 << 24, (n,), device="cuda", dtype=torch.int64, generator=g)
    elif dt == torch.int64:
        k0 = torch.randint(0, 1 << 62, (n,), device="cuda", dtype=torch.int64, generator=g) * 2 + \
             torch.randint(0, 2, (n,), device="cuda", dtype=torch.int64, generator=g)
    else:
        k0 = torch.randint(0, 1 << 30, (n,), device="cuda", dtype=torch.int32, generator=g)
    srt = ops.SfcSorter("cuda")
    keys = k0.clone()
    order = srt.set_map_from_codes(keys)
    ref_k, ref_o = torch.sort(k0, stable=True)
    ok = bool(torch.equal(keys, ref_k)) and bool(torch.equal(order.long(), ref_o))
    for iota in (True, False):
        ms = []
        for _ in range(reps):
            keys.copy_(k0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if iota: srt.set_map_from_codes(keys)
            else: srt.update_map(keys)
            e1.record(); torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        ms.sort(); med = ms[len(ms) // 2]
        esz = k0.element_size()
        alg = n * (esz + passes * 2 * (esz + 4) - (4 if iota else 0))
        print("%s n=%d iota=%d parity=%s  %.3f ms (min %.3f)  %.0f GB/s algorithmic" %
              (name, n, iota, ok, med, ms[0], alg / med / 1e6))
