"""Device timings of the focus-tree / expansion-centre / target-group kernels at hot-path size (not part of bench.py's
step; the numbers go into DESIGN.md §4).  CUDA events on the current stream, warm-up + N repetitions, L2 flushed
between repetitions by a 256 MB memset.  Usage: python tools/bench_focus.py [n_particles] [--json out.json]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cornerstone_octree_b200 import ops  # noqa: E402
from cornerstone_octree_b200.focus_tree import FocusedOctreeSingleNode  # noqa: E402


def timed(fn, flush, reps=5, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 32_000_000
    out = sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None
    dev = "cuda"
    peak = 6550.1
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    g = torch.Generator(device=dev).manual_seed(1)
    x, y, z = (torch.rand(n, dtype=torch.float64, device=dev, generator=g) for _ in range(3))
    box = ops.make_box((0.0, 1.0), dtype=torch.float64)
    keys = ops.compute_sfc_keys(x, y, z, box)
    keys, order = ops.sort_by_key(keys)
    x, y, z = (ops.gather(order, a) for a in (x, y, z))
    m = torch.rand(n, dtype=torch.float64, device=dev, generator=g) + 0.5
    h = torch.full((n,), 0.005, dtype=torch.float64, device=dev)
    la = ops.LeafArray(64, capacity=max(1024, n // 8))
    la.compute(keys, 64)
    leaves, counts = la.tree.clone(), la.counts.clone()
    L = leaves.numel() - 1
    oc = ops.build_octree(leaves)
    M = oc.num_nodes
    layout = torch.zeros(L + 1, dtype=torch.int32, device=dev)
    layout[1:] = torch.cumsum(counts.to(torch.int64), 0).to(torch.int32)
    lr = oc.level_range.cpu().numpy()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    res = dict(n_particles=n, num_leaves=L, num_nodes=M, hbm_peak_gbps=peak, kernels={})

    def rec(name, ms, nbytes):
        gbps = nbytes / (ms * 1e-3) / 1e9
        res["kernels"][name] = dict(ms=round(ms, 4), algorithmic_bytes=int(nbytes), gbps=round(gbps, 1),
                                    frac_of_hbm_peak=round(gbps / peak, 3))
        print("%-34s %9.3f ms  %8.1f GB/s  %5.1f %% of HBM peak" % (name, ms, gbps, 100 * gbps / peak), flush=True)

    # expansion centres
    c4 = torch.zeros((M, 4), dtype=torch.float64, device=dev)
    rec("compute_leaf_source_center", timed(lambda: ops.compute_leaf_source_center(x, y, z, m, oc, layout, c4), flush),
        32 * n + 32 * L + 8 * L)
    rec("upsweep_centers", timed(lambda: ops.upsweep_centers(lr, oc.child_offsets, c4), flush), 32 * M + 32 * (M - L) + 4 * M)
    # groups
    for gs in (32, 64):
        rec("compute_group_splits_%d" % gs,
            timed(lambda: ops.compute_group_splits(0, n, x, y, z, h, leaves, layout, box, gs, 2.0), flush),
            24 * n + 4 * (n // gs) * 3)
    ng = ops.compute_group_splits(0, n, x, y, z, h, leaves, layout, box, 64, 2.0).numel() - 1
    res["groups_64_tol2"] = int(ng)
    # node counts for the decisions
    nc = torch.zeros(M, dtype=torch.int32, device=dev)
    nc[oc.leaf_to_internal[oc.num_internal_nodes:].long()] = counts
    ops.upsweep_sum(lr, oc.child_offsets, nc)
    # MAC spheres + marks for a focus = 1/8 of the leaves (one rank of 8)
    f0, f1 = 3 * L // 8, 4 * L // 8
    inv_theta = 1.0 / 0.5 + 0.5
    rec("geo_mac_spheres", timed(lambda: ops.geo_mac_spheres(oc.prefixes, inv_theta, box), flush), 8 * M + 32 * M)
    cen = ops.geo_mac_spheres(oc.prefixes, inv_theta, box)
    rec("set_mac", timed(lambda: ops.set_mac(oc.prefixes, cen, inv_theta, box), flush), 8 * M + 64 * M)
    cen = ops.geo_mac_spheres(oc.prefixes, inv_theta, box)
    focus = leaves[f0:f1 + 1].contiguous()
    macs = torch.zeros(M, dtype=torch.int8, device=dev)

    def marks():
        macs.zero_()
        ops.mark_macs(oc, cen, box, focus, False, macs)

    ms = timed(marks, flush)
    res["mark_macs"] = dict(ms=round(ms, 4), focus_leaves=int(f1 - f0), marked=int(macs.sum().item()),
                            ns_per_focus_leaf=round(ms * 1e6 / (f1 - f0), 2))
    print("%-34s %9.3f ms  (%d focus leaves, %d nodes marked)" % ("mark_macs", ms, f1 - f0, res["mark_macs"]["marked"]),
          flush=True)
    fs, fe = int(leaves[f0].item()), int(leaves[f1].item())
    rec("rebalance_decision_essential",
        timed(lambda: ops.rebalance_decision_essential(oc, nc, macs, fs, fe, 64), flush), (8 + 4 + 4 + 1 + 4) * M)
    nops = ops.rebalance_decision_essential(oc, nc, macs, fs, fe, 64)
    rec("protect_ancestors", timed(lambda: ops.protect_ancestors(oc, nops.clone()), flush), (8 + 4 + 4 + 4) * M)
    rec("mac_refine_decision", timed(lambda: ops.mac_refine_decision(oc, macs, f0, f1), flush), (4 + 8 + 1 + 4) * L)
    # range count: focus tree = the same leaves coarsened to bucket 512 (global -> focus counts)
    lb = ops.LeafArray(64, capacity=max(1024, n // 64))
    lb.compute(keys, 512)
    ftree = lb.tree.clone()
    idx = torch.arange(ftree.numel() - 1, dtype=torch.int32, device=dev)
    rec("range_count", timed(lambda: ops.range_count(leaves, counts, ftree, idx), flush), 12 * L + 16 * idx.numel())

    # whole focus-tree update: steps to convergence from the root and the steady-state step
    ft = FocusedOctreeSingleNode(64, 0.5)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    steps = 1
    while not ft.update(box, keys, fs, fe):
        steps += 1
        assert steps < 64
    e1.record()
    torch.cuda.synchronize()
    build_ms = e0.elapsed_time(e1)
    ms = timed(lambda: ft.update(box, keys, fs, fe), flush, reps=5, warm=1)
    fl = ft.tree_leaves().numel() - 1
    res["focus_tree"] = dict(focus_leaves=int(fl), build_steps=steps, build_ms=round(build_ms, 3),
                             steady_update_ms=round(ms, 3))
    print("focus tree: %d leaves (focus = 1/8 of the curve, theta 0.5), %d steps / %.2f ms from the root, %.3f ms per "
          "steady-state update" % (fl, steps, build_ms, ms), flush=True)
    if out:
        json.dump(res, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
