#!/usr/bin/env python
"""bench.py — Domain::sync + findNeighbors throughput (Mparticles/s) of the B200 hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the whole hot path over one rank's particles:
  computeSfcKeys -> sort (key + permutation) -> gather x,y,z,h -> cornerstone leaf array (built from the root)
  -> internal-tree linking -> node centres -> layout -> halo radii -> findHalos (8 virtual ranks, every leaf searched
  once as a local leaf) -> findNeighbors(ngmax) over all particles.
Workload at N=1: BASELINE.json configs[1] — 32 M uniform particles, u64 Hilbert keys, double, bucket 64 — with the
metric's findNeighbors (ngmax 150, ~100 neighbours per particle) appended.

The JSON line follows the contract in the task text: `value` is device-resident throughput, `e2e` goes through the
C ABI from pinned HOST buffers (H2D of x,y,z,h and D2H of the results inside the timed region), `roofline` is the
dominant kernel against MEASURED_PEAKS.json, `cpu_baseline` is the reference CPU code (oracle/_ref) timed on the
host cores on a bounded sample.  `--impl reference` times only that CPU code (rank 0).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "Domain::sync + findNeighbors"
UNIT = "Mparticles/s"
NG0 = 100


# ------------------------------------------------------------------------------------------------ inputs
def make_particles(n, dist, seed):
    """seeded synthetic cloud on the host: x,y,z,h (float64) and the box limits"""
    rng = np.random.default_rng(seed)
    if dist == "uniform":
        x, y, z = rng.random(n), rng.random(n), rng.random(n)  # all x, then all y, then all z
        lim = [0.0, 1.0] * 3
        h = np.full(n, 0.5 * (3.0 * NG0 / (4.0 * math.pi * n)) ** (1.0 / 3.0))
    elif dist == "plummer":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import gen
        (x, y, z), lim = gen.plummer(n, np.float64, seed)
        # local-density smoothing length: Plummer density profile, ~NG0 neighbours inside 2h
        a = 3.0 * math.pi / 16.0
        r2 = x * x + y * y + z * z
        rho = n * 3.0 / (4.0 * math.pi * a ** 3) * (1.0 + r2 / (a * a)) ** -2.5
        h = 0.5 * (3.0 * NG0 / (4.0 * math.pi * rho)) ** (1.0 / 3.0)
        h = np.minimum(h, 0.05 * (lim[1] - lim[0]))
    else:
        raise ValueError(dist)
    return x, y, z, h, lim


# ------------------------------------------------------------------------------------------------ CPU reference
def cpu_pipeline(orc, x, y, z, h, box, bucket, ngmax, vranks=8):
    """the same pass through the reference's CPU code (oracle/_ref) or our CPU port of it; returns mean neighbours"""
    kind = 1
    keys = orc.compute_sfc_keys(kind, x, y, z, box, np.uint64)
    skeys, order = orc.sort_by_key(keys)
    xs, ys, zs, hs = (orc.gather(order, a) for a in (x, y, z, h))
    tree, counts = orc.compute_octree(skeys, bucket)
    oc = orc.build_octree(tree)
    centers, sizes = orc.geo_centers(kind, oc["prefixes"], box, np.float64)
    L = tree.size - 1
    layout = np.zeros(L + 1, dtype=np.uint32)
    np.cumsum(counts, out=layout[1:])
    radii = orc.halo_radii(hs, layout, 2.0)
    for r in range(vranks):
        first, last = r * L // vranks, (r + 1) * L // vranks
        rr = np.zeros_like(radii)
        rr[first:last] = radii[first:last]
        orc.find_halos(kind, oc, tree, rr, box, first, last, np.float64)
    nb, nc = orc.find_neighbors(xs, ys, zs, hs, 0, x.size, box, oc, layout, centers, sizes, ngmax)
    return float(nc.mean())


def time_cpu(n, dist, bucket, ngmax, steps, warmup):
    from oracle import loader
    which = "ref" if loader.available("ref") else "port"
    if which == "port" and not loader.available("port"):
        loader.build("port")
    orc = loader.load(which)
    cores = os.cpu_count() or 1
    orc.set_num_threads(cores)
    x, y, z, h, lim = make_particles(n, dist, 42)
    box = loader.make_box(lim)
    for _ in range(warmup):
        cpu_pipeline(orc, x, y, z, h, box, bucket, ngmax)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_pipeline(orc, x, y, z, h, box, bucket, ngmax)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    kind = "reference" if which == "ref" else "port"
    return dict(value=n / dt / 1e6, unit=UNIT, cores=cores, kind=kind,
                sample="%d %s particles, same pipeline, oracle/%s, OpenMP %d threads, %.2f s/step" %
                       (n, dist, "_ref" if which == "ref" else "port", cores, dt)), dt


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.path = "/tmp/cstone_clocks_%d_%d.csv" % (os.getpid(), index)
        self.proc = None
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[])
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            p = [s.strip() for s in line.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if p[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------ main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=32_000_000, help="particles per GPU")
    ap.add_argument("--dist", default="uniform", choices=["uniform", "plummer"])
    ap.add_argument("--bucket", type=int, default=64)
    ap.add_argument("--ngmax", type=int, default=150)
    ap.add_argument("--cpu-n", type=int, default=2_000_000, help="particles of the bounded CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--steady", action="store_true",
                    help="steady-state sync: one leaf-array update step from the previous tree instead of a build "
                         "from the root")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    workload = ("%dM %s particles per GPU, u64 Hilbert keys, double x/y/z/h, bucketSize=%d: key->sort->gather->"
                "leaf array (%s)->link->centres->layout->radii->findHalos(8 virtual ranks)->findNeighbors(ngmax=%d)" %
                (args.n // 1_000_000, args.dist, args.bucket,
                 "update step" if args.steady else "computeOctree from root", args.ngmax))

    if args.impl == "reference":
        if rank != 0:
            return 0
        base, dt = time_cpu(args.cpu_n, args.dist, args.bucket, args.ngmax, args.steps, args.warmup)
        line = dict(metric=METRIC, value=base["value"], unit=UNIT, n_gpus=args.gpus, steps=args.steps,
                    warmup=args.warmup, ms_per_step=dt * 1e3, higher_is_better=True, scaling="weak",
                    vs_baseline=None, dtype="f64", data="synthetic", impl="reference",
                    config=dict(workload=workload, sample=base["sample"]), cpu_baseline=base,
                    e2e=dict(value=base["value"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                    gpu_launches=0)
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import cornerstone_octree_b200 as pkg
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200.pipeline import HotPath

    n = args.n
    x, y, z, h, lim = make_particles(n, args.dist, 42 + rank)
    box = ops.make_box(lim)
    host = [torch.from_numpy(a).pin_memory() for a in (x, y, z, h)]
    dev = [t.cuda(non_blocking=True) for t in host]
    torch.cuda.synchronize()
    hp = HotPath(n, 64, torch.float64, args.bucket, args.ngmax)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cold = not args.steady

    steady_in = []

    def run_step(time_stages=False):
        if args.steady and steady_in:
            return hp.step(*steady_in, box, cold=False, time_stages=time_stages)
        Lx = hp.step(*dev, box, cold=True, time_stages=time_stages)
        if args.steady:  # later steps start from the SFC-ordered arrays and the previous leaf array
            steady_in.extend(t.clone() for t in hp.sorted)
        return Lx

    # ---- warm-up (also sizes the leaf / tree buffers)
    for _ in range(max(args.warmup, 1)):
        L = run_step()
    barrier()

    # ---- device-resident timing: EXACTLY K steps between two events
    sampler = ClockSampler(local_rank)
    launches0 = pkg.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_acc = {}
    barrier()
    ev0.record()
    stage_events = []
    for _ in range(args.steps):
        L = run_step(time_stages=True)
        stage_events.append(hp.events)
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    launches = pkg.launch_count() - launches0
    clocks = sampler.stop()
    for evs in stage_events:
        for (_, a), (name, b) in zip(evs[:-1], evs[1:]):
            stage_acc[name] = stage_acc.get(name, 0.0) + a.elapsed_time(b) / args.steps
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = n * world / (ms_step * 1e-3) / 1e6

    nc_mean = float(hp.neighbors_count.to(torch.float64).mean().item())
    nc_stored = float(torch.clamp(hp.neighbors_count, max=args.ngmax).to(torch.float64).mean().item())

    # ---- end to end from pinned host buffers through the same calls
    e2e = None
    if not args.no_e2e:
        M = hp.octree.num_nodes
        out_host = dict(keys=torch.empty(n, dtype=torch.int64).pin_memory(),
                        nc=torch.empty(n, dtype=torch.int32).pin_memory(),
                        tree=torch.empty(L + 1, dtype=torch.int64).pin_memory(),
                        counts=torch.empty(L, dtype=torch.int32).pin_memory(),
                        flags=torch.empty((hp.vranks, L), dtype=torch.int32).pin_memory(),
                        sorted=[torch.empty(n, dtype=torch.float64).pin_memory() for _ in range(4)])
        dstage = [torch.empty_like(d) for d in dev]

        def e2e_step():
            for d, s in zip(dstage, host):
                d.copy_(s, non_blocking=True)
            Lx = hp.step(*dstage, box, cold=True)
            out_host["keys"].copy_(hp.keys, non_blocking=True)
            out_host["nc"].copy_(hp.neighbors_count, non_blocking=True)
            out_host["tree"][:Lx + 1].copy_(hp.leaves.tree_a[:Lx + 1], non_blocking=True)
            out_host["counts"][:Lx].copy_(hp.leaves.counts_buf[:Lx], non_blocking=True)
            for r in range(hp.vranks):
                out_host["flags"][r, :Lx].copy_(hp.flags[r, :Lx], non_blocking=True)
            for o, s in zip(out_host["sorted"], hp.sorted):
                o.copy_(s, non_blocking=True)
            torch.cuda.synchronize()
            return int(out_host["nc"][0])

        e2e_step()
        barrier()
        ev0.record()
        for _ in range(args.steps):
            e2e_step()
        ev1.record()
        barrier()
        te = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        ms_e2e = float(te.item()) / args.steps
        h2d = 4 * n * 8
        d2h = n * 8 + n * 4 + (L + 1) * 8 + L * 4 + hp.vranks * L * 4 + 4 * n * 8
        e2e = dict(value=n * world / (ms_e2e * 1e-3) / 1e6, unit=UNIT, h2d_bytes_per_step=h2d,
                   d2h_bytes_per_step=d2h, ms_per_step=ms_e2e,
                   outputs="sorted keys + x,y,z,h, leaf array + counts, halo flags, neighbour counts; the neighbour "
                           "lists (%d x %d x 4 B) stay device-resident as for Domain<...,GpuTag> clients" %
                           (n, args.ngmax))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel + per-stage table
    peak, peak_src = load_peaks()
    sK, sT = 8, 8
    alg = {
        "sfc_keys": n * (3 * sT + 2 * sK),
        "sort": n * (sK + 8 * 2 * (sK + 4) - 4),          # histogram read + 8 passes r/w of (key, u32); pass 0 reads no value
        "gather": n * (4 + 4 * 2 * sT),
        "find_neighbors": n * (4 * sT + 4) + 4.0 * nc_stored * n,
    }
    stages = {}
    for name, ms in stage_acc.items():
        d = dict(ms=round(ms, 4), share=round(ms / ms_step, 4))
        if name in alg:
            gbs = alg[name] / (ms * 1e-3) / 1e9
            d.update(alg_bytes=int(alg[name]), gbs=round(gbs, 1), frac=round(gbs / peak, 4))
        stages[name] = d
    dom = max(("find_neighbors", "sort", "sfc_keys", "gather"), key=lambda k: stage_acc.get(k, 0.0))
    kernel_name = {"find_neighbors": "findNeighborsKernel<double>", "sort": "onesweepKernel<u64> x8 + histogram",
                   "sfc_keys": "sfcKeysKernel", "gather": "gatherMultiKernel"}[dom]
    ach = alg[dom] / (stage_acc[dom] * 1e-3) / 1e9
    roofline = dict(bound="hbm", kernel=kernel_name, achieved=round(ach, 1), peak=peak, unit="GB/s",
                    frac=round(ach / peak, 4), traffic=None, peak_source=peak_src,
                    alg_bytes_per_launch=int(alg[dom]), launch_ms=round(stage_acc[dom], 4))

    cpu = None
    if not args.no_cpu and world == 1:
        cpu, _ = time_cpu(args.cpu_n, args.dist, args.bucket, args.ngmax, 1, 0)

    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 1),
                ms_per_step=ms_step, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64",
                data="synthetic", impl="b200",
                config=dict(workload=workload, particles_per_gpu=n, leaves=int(L), neighbours_mean=round(nc_mean, 2),
                            l2="inputs (>=1 GB per stage) exceed the 126 MB L2; no explicit flush",
                            parallelism="1 rank per GPU" if world == 1 else "%d independent SFC shards" % world),
                clocks=clocks, e2e=e2e, gpu_launches=int(launches), roofline=roofline, stages=stages,
                cpu_baseline=cpu)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
