#!/usr/bin/env python
"""bench.py — Domain::sync + findNeighbors throughput (Mparticles/s) of the B200 hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one steady-state pass of the hot path over every rank's particles: a small deterministic drift, then
  Domain::sync (box, keys, sort, global tree + assignment, particle exchange, halo discovery + exchange, tree update
  from the previous step's tree, linking, centres) and findNeighbors(ngmax) over the assigned particles
  (+ findHalos for 8 virtual ranks at N = 1, BASELINE configs[1]).
Workload at N=1: BASELINE.json configs[1] — 32 M uniform particles, u64 Hilbert keys, double, bucket 64 — with the
metric's findNeighbors (ngmax 150, ~100 neighbours per particle) appended.  --fields 4 adds the four conserved
double fields of configs[3]; --dist plummer is configs[2].

The JSON line follows the contract in the task text: `value` is device-resident throughput, `e2e` goes through the
C ABI from pinned HOST buffers (H2D of x,y,z,h and D2H of the results inside the timed region), `roofline` is the
dominant kernel against MEASURED_PEAKS.json, `check` is correctness evidence gathered inside this run.
Reference arm (`--impl reference`, and the `cpu_baseline` leg of the default run at N = 1): the UNMODIFIED reference
class Domain<KeyType,T,CpuTag>::sync + findNeighbors (oracle/_ref/liboracle_refdom.so: reference headers over a
thread-rank mpi.h) on the host cores, one thread-rank per GPU of the run, the same generator, drift and steady-state
stepping; at most --ref-total particles in total (32 M: the whole N = 1 workload).
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
# stdout carries exactly one JSON line: NCCL's own banner / debug output (NCCL_DEBUG=VERSION|INFO in the environment)
# goes to stderr.  Must be set before the first communicator is created.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":  # NCCL honours NCCL_DEBUG_FILE only above level VERSION
    os.environ["NCCL_DEBUG"] = "WARN"

METRIC = "Domain::sync + findNeighbors"
UNIT = "Mparticles/s"
NG0 = 100
# dram__bytes_read.sum + dram__bytes_write.sum of one findNeighborsKernel launch at 32 M uniform (ncu --set full)
NEIGHBOR_TRAFFIC_BYTES = 23_385_864_000  # dram read (8.688 GB) + write (14.698 GB) of one launch, profiles/r01g_findNeighbors_full.txt


# ------------------------------------------------------------------------------------------------ inputs
def mt19937_uniform(n, lo, hi, bg):
    """n draws of std::uniform_real_distribution<double>(lo, hi) from a std::mt19937 (libstdc++): generate_canonical
    takes two 32-bit outputs g1, g2 and returns (g1 + g2 * 2^32) / 2^64 in double arithmetic"""
    raw = bg.random_raw(2 * n).astype(np.float64)
    u = (raw[0::2] + raw[1::2] * 4294967296.0) / 18446744073709551616.0
    u = np.where(u >= 1.0, np.nextafter(1.0, 0.0), u)
    return u * (hi - lo) + lo


def make_particles(n, dist, seed):
    """seeded synthetic cloud on the host: x,y,z,h (float64) and the box limits"""
    if dist == "uniform":
        # SURVEY 8d / test/coord_samples/random.hpp:102-113: std::mt19937 gen(seed), all x, then all y, then all z
        bg = np.random.MT19937()
        bg._legacy_seeding(seed)
        x, y, z = (mt19937_uniform(n, 0.0, 1.0, bg) for _ in range(3))
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
def drift_amplitude(h0):
    return float(0.1 * h0)


def time_reference(args, world, steps, warmup, total_cap, full_n=None):
    """the reference's Domain<uint64_t,double,CpuTag>::sync + findNeighbors on `world` thread-ranks of this process
    (oracle/_ref/liboracle_refdom.so), stepped exactly like the GPU arm: cold first sync, `warmup` steady steps, then
    `steps` timed ones.  Returns (cpu_baseline dict, seconds per step, RefDomain or None)"""
    from oracle import loader
    if not loader.refdom_available():
        raise SystemExit("bench.py: oracle/_ref/liboracle_refdom.so is missing (run __graft_entry__.build() where the "
                         "reference tree is present)")
    cores = os.cpu_count() or 1
    full_n = full_n or args.n
    n_rank = min(args.n, max(total_cap // world, 1))
    lim_all, slabs = None, []
    for r in range(world):
        x, y, z, h, lim = make_particles(n_rank, args.dist, 42 + r)
        if args.dist == "uniform" and world > 1:
            h = h / world ** (1.0 / 3.0)  # same ~100 neighbours at the global density (as the GPU arm)
        slabs.append((x, y, z, h))
        lim_all = list(lim) if lim_all is None else [f(a, b) for f, a, b in zip([min, max] * 3, lim_all, lim)]
    nprops = args.fields if args.fields in (0, 1, 2, 4) else 0
    rd = loader.RefDomain(world, args.global_bucket, args.bucket, 0.5, lim_all, nprops=nprops,
                          kdtype=np.uint64 if args.keys == "u64" else np.uint32, omp_per_rank=max(1, cores // world))
    for r, (x, y, z, h) in enumerate(slabs):
        gid = np.arange(r * n_rank, (r + 1) * n_rank, dtype=np.float64)
        rd.set(r, x, y, z, h, props=[gid, 2 * gid][: min(nprops, 2)])  # 4 fields: id * {1,2,3,4}, derived inside
    amp = drift_amplitude(slabs[0][3][0])
    del slabs
    sign = 1.0
    t_first = time.perf_counter()
    rd.bench_step(amp * sign, args.ngmax)  # cold: distributes the unsorted slabs, builds the trees from the root
    first_s = time.perf_counter() - t_first
    sign = -sign
    for _ in range(warmup):
        rd.bench_step(amp * sign, args.ngmax)
        sign = -sign
    t0 = time.perf_counter()
    tot = (0, 0.0, 0)
    for _ in range(steps):
        tot = rd.bench_step(amp * sign, args.ngmax)
        sign = -sign
    dt = (time.perf_counter() - t0) / max(steps, 1)
    assigned, nc_sum, halos = tot
    sync_s, nb_s = rd.seconds()
    base = dict(value=assigned / dt / 1e6, unit=UNIT, cores=cores, kind="reference",
                sample="%d thread-rank(s) x %d %s particles = %d in total%s: reference Domain<%s,double,CpuTag>::sync + "
                       "findNeighbors(ngmax %d), steady state, OpenMP %d thread(s) per rank; %.2f s/step (sync %.2f, "
                       "findNeighbors %.2f), first sync %.1f s, %.1f neighbours / particle" %
                       (world, n_rank, args.dist, assigned,
                        "" if n_rank == full_n else " (bounded: the GPU arm runs %d per rank)" % full_n, args.keys, args.ngmax,
                        max(1, cores // world), dt, sync_s, nb_s, first_s, nc_sum / max(assigned, 1)),
                particles=assigned, same_size=bool(n_rank == full_n))
    return base, dt, rd


def parity_in_run(args, n_check=1_000_000):
    """correctness inside the run (N = 1): a fresh GPU Domain and the reference Domain on the same n_check-particle
    cloud, one cold sync + findNeighbors each; keys, coordinates, tree leaves, neighbour counts and the raw neighbour
    lists must be identical"""
    import torch
    from oracle import loader
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200.domain import Domain
    n = min(n_check, args.n)
    if args.curve != "hilbert":
        return dict(particles=0, ok=True, skipped="the reference's Domain hard-wires the Hilbert curve")
    kd = np.uint64 if args.keys == "u64" else np.uint32
    x, y, z, h, lim = make_particles(n, args.dist, 4242)
    rd = loader.RefDomain(1, args.global_bucket, args.bucket, 0.5, lim, kdtype=kd)
    rd.set(0, x, y, z, h)
    rd.bench_step(0.0, args.ngmax)
    rk, rx, ry, rz, rh = rd.arrays(0)[:5]
    rnb, rnc = rd.bench_neighbors(0, args.ngmax)
    rleaves = rd.trees(0)["focus_leaves"]
    rd.close()
    dom = Domain(0, 1, args.global_bucket, args.bucket, 0.5, ops.make_box(lim), key_bits=8 * np.dtype(kd).itemsize)
    keys, sx, sy, sz, sh, _ = dom.sync(*[torch.from_numpy(a).cuda() for a in (x, y, z, h)])
    nb, nc = dom.find_neighbors(sx, sy, sz, sh, args.ngmax)
    leaves = dom.tree_arrays()["leaves"].cpu().numpy().view(kd)
    nb, nc = nb.cpu().numpy().view(np.uint32), nc.cpu().numpy().view(np.uint32)
    stored = np.arange(args.ngmax)[None, :] < np.minimum(nc, args.ngmax)[:, None]
    res = dict(particles=int(n),
               keys=bool(np.array_equal(keys.cpu().numpy().view(kd), rk)),
               coordinates=bool(all(np.array_equal(a.cpu().numpy(), b) for a, b in zip((sx, sy, sz, sh), (rx, ry, rz, rh)))),
               leaves=bool(np.array_equal(leaves, rleaves)),
               neighbour_counts=bool(np.array_equal(nc, rnc)),
               neighbour_lists=bool(np.array_equal(np.where(stored, nb, 0), np.where(stored, rnb, 0))))
    dom.close()
    res["ok"] = all(v for k, v in res.items() if k != "particles")
    return res


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi sampling of SM clocks / throttle reasons.  Started BEFORE the steady warm-up steps (nvidia-smi needs
    up to a second to come up on an 8-GPU box, longer than a short timed region); every sample carries a timestamp and
    stop(t0, t1) keeps the samples inside the timed region [t0, t1] (host clock).  If the region was too short to catch
    one, the samples of the steady warm-up steps (same load) are used and `window` says so."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,timestamp")

    def __init__(self, index):
        self.path = "/tmp/cstone_clocks_%d_%d.csv" % (os.getpid(), index)
        self.proc = None
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=self.f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self, t0=None, t1=None):
        import datetime
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[])
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []
        for line in open(self.path):
            p = [s.strip() for s in line.split(",")]
            if len(p) < 8:
                continue
            try:
                row = (float(p[0]), float(p[1]), [nm for k, nm in enumerate(names) if p[3 + k].lower().startswith("active")])
            except ValueError:
                continue
            try:
                ts = datetime.datetime.strptime(p[7], "%Y/%m/%d %H:%M:%S.%f").timestamp()
            except ValueError:
                ts = None
            rows.append((ts,) + row)
        inside = [r for r in rows if t0 is not None and r[0] is not None and t0 <= r[0] <= t1]
        out["window"] = "timed region" if inside else "warm-up + timed region"
        if not inside and t0 is not None:  # prefer the samples closest to the timed region (steady warm-up steps)
            late = [r for r in rows if r[0] is not None and r[0] >= t0 - 1.0]
            rows = late or rows
        sm, mx, reasons = [], [], set()
        for _, a, b, rs in (inside or rows):
            sm.append(a)
            mx.append(b)
            reasons.update(rs)
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
    ap.add_argument("--n", "--particles", dest="n", type=int, default=32_000_000, help="particles per GPU")
    ap.add_argument("--dist", default="uniform", choices=["uniform", "plummer"])
    ap.add_argument("--bucket", type=int, default=64)
    ap.add_argument("--ngmax", type=int, default=150)
    ap.add_argument("--keys", default="u64", choices=["u64", "u32"], help="SFC key width (configs[4]: u32)")
    ap.add_argument("--curve", default="hilbert", choices=["hilbert", "morton"],
                    help="space-filling curve (configs[4]: morton; the reference's Domain hard-wires Hilbert)")
    ap.add_argument("--fields", type=int, default=0, help="conserved double fields carried through sync (configs[3]: 4)")
    ap.add_argument("--ref-total", type=int, default=32_000_000,
                    help="upper bound of the particles (all thread-ranks together) of the reference arm")
    ap.add_argument("--cpu-n", type=int, default=8_000_000, help="particles of the bounded cpu_baseline leg (N = 1)")
    ap.add_argument("--halo-mode", type=int, default=0, choices=[0, 1],
                    help="N > 1: 0 = fast halo discovery (default), 1 = the reference's procedure on the multi-rank focus tree")
    ap.add_argument("--halo-push", type=int, default=1, choices=[0, 1],
                    help="N > 1: 1 = pack-and-push halo exchange over peer memory (default), 0 = gather + grouped send/recv")
    ap.add_argument("--no-check", action="store_true", help="skip the correctness checks inside the run")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--global-bucket", type=int, default=65536, help="bucketSize of the replicated global tree")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    nranks = args.gpus if args.impl == "reference" else world
    bits = 64 if args.keys == "u64" else 32
    keyname = "%s %s" % (args.keys, args.curve.capitalize())
    workload = ("%dM %s particles per GPU, %s keys, double x/y/z/h%s, bucketSizeFocus=%d: steady-state "
                "Domain::sync (box, keys, sort, global tree + assignment, particle exchange, halo discovery + exchange, "
                "tree update, linking, centres) + findNeighbors(ngmax=%d) over the assigned particles%s" %
                (args.n // 1_000_000, args.dist, keyname, " + %d double fields" % args.fields if args.fields else "",
                 args.bucket, args.ngmax, "; + findHalos for 8 virtual ranks (configs[1])" if nranks == 1 else ""))
    # the same dictionary in both arms: what is computed, on what
    config = dict(workload=workload, particles_per_gpu=args.n, ranks=nranks, dist=args.dist, keys=keyname,
                  real="f64", bucket_focus=args.bucket, bucket_global=args.global_bucket, ngmax=args.ngmax,
                  fields=args.fields,
                  generator="std::mt19937(42 + rank) + uniform_real_distribution, all x then y then z (SURVEY 8d); "
                            "h = 0.5 cbrt(3 * 100 / (4 pi n_global))" if args.dist == "uniform" else
                            "tests/gen.py plummer(seed 42 + rank), h from the local Plummer density",
                  drift="0.1 h deterministic displacement before every sync, inside the timed region",
                  l2="inputs (>= 1 GB per stage) exceed the 126 MB L2; no explicit flush")

    if args.impl == "reference":
        # rank 0 alone runs the reference on the host cores, one thread-rank per GPU of the run; the other ranks leave
        if rank != 0:
            return 0
        base, dt, rd = time_reference(args, args.gpus, args.steps, args.warmup, args.ref_total)
        rd.close()
        if not base["same_size"]:
            config = dict(config, reference_sample="%d particles per thread-rank instead of %d (bounded run time)" %
                          (base["particles"] // args.gpus, args.n))
        line = dict(metric=METRIC, value=base["value"], unit=UNIT, n_gpus=args.gpus, steps=args.steps,
                    warmup=args.warmup, ms_per_step=dt * 1e3, higher_is_better=True, scaling="weak",
                    vs_baseline=None, dtype="f64", data="synthetic", impl="reference", config=config,
                    cpu_baseline=base,
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
    import ctypes as C
    import cornerstone_octree_b200 as pkg
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200._lib import call
    from cornerstone_octree_b200.domain import Domain

    n = args.n
    # every rank starts with its slab of an UNSORTED global cloud (as test/integration_mpi/domain_nranks.cpp:103-109)
    x, y, z, h, lim = make_particles(n, args.dist, 42 + rank)
    if args.dist == "uniform" and world > 1:
        h = h / world ** (1.0 / 3.0)  # same ~100 neighbours at the global density of world * n particles
    box = ops.make_box(lim)
    sfc_kind = ops.HILBERT if args.curve == "hilbert" else ops.MORTON
    dom = Domain(rank, world, args.global_bucket, args.bucket, 0.5, box, key_bits=bits, sfc_kind=sfc_kind)
    if world > 1 and args.halo_mode:
        dom.set_halo_mode(args.halo_mode)
    if world > 1:
        dom.set_halo_push(bool(args.halo_push))
    dom.init_comm_from_torch()
    dev = [torch.from_numpy(a).cuda() for a in (x, y, z, h)]
    del x, y, z
    amp = drift_amplitude(h[0])
    # conserved fields (configs[3]): double(global id) * {1, 2, ...}, so that what arrives can be checked
    gid = torch.arange(rank * n, (rank + 1) * n, dtype=torch.float64, device="cuda")
    fields = [gid * float(k + 1) for k in range(args.fields)]
    del gid
    state = {"a": dev, "f": fields, "sign": 1.0, "nb": None, "nc": None, "flags": None}
    vranks = 8

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def drift(arrs):
        """small deterministic displacement (0.1 h) so that every sync sees moved particles"""
        xx, yy, zz, hh = arrs
        w = 2.0 * math.pi * 37.0
        xx.add_(torch.sin(yy * w), alpha=amp * state["sign"])
        yy.add_(torch.sin(zz * w), alpha=amp * state["sign"])
        zz.add_(torch.sin(xx * w), alpha=amp * state["sign"])
        state["sign"] = -state["sign"]

    def halos_single_rank(t_leaves, n_leaves, info):
        """N = 1 only: a single rank has no halos, so the halo search of BASELINE configs[1] is run for 8 virtual
        ranks on the Domain's tree: every leaf is searched exactly once as a local leaf"""
        ptrs = info
        if state["flags"] is None or state["flags"].shape[1] < n_leaves:
            state["flags"] = torch.empty((vranks, int(n_leaves * 1.1) + 64), dtype=torch.int32, device="cuda")
            state["radii"] = torch.empty(int(n_leaves * 1.1) + 64, dtype=torch.float32, device="cuda")
        fl = state["flags"]
        fl.zero_()
        call("halo_radii_f64", ops._p(state["a"][3]), C.c_void_p(ptrs[2]), n_leaves, 2.0, ops._p(state["radii"]), ops._stream())
        bx = dom.box()
        # the 8 searches are independent (own flag arrays): one stream each, so that a 262 k-leaf launch (less than
        # one wave of threads) does not leave the GPU idle; the main stream waits for all of them
        if state.get("hstreams") is None:
            state["hstreams"] = [torch.cuda.Stream() for _ in range(vranks)]
        main = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(main)
        for r in range(vranks):
            first, last = r * n_leaves // vranks, (r + 1) * n_leaves // vranks
            st = state["hstreams"][r]
            st.wait_event(ready)
            call("find_halos_k%d_f64" % bits, int(sfc_kind), C.c_void_p(ptrs[3]), C.c_void_p(ptrs[4]), C.c_void_p(ptrs[5]), C.c_void_p(ptrs[0]),
                 ops._p(state["radii"]), C.byref(bx), first, last, ops._p(fl[r]), C.c_void_p(st.cuda_stream))
            done = torch.cuda.Event()
            done.record(st)
            main.wait_event(done)

    from cornerstone_octree_b200.domain import _fn, _check
    tree_ptrs = (C.c_void_p * 10)()

    def run_step(marks=None, after_sync=None):
        def mark(name):
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append((name, e))
        mark("start")
        drift(state["a"])
        mark("drift")
        keys, sx, sy, sz, sh, fl = dom.sync(*state["a"], props=state["f"], fields_ready=state.pop("h_ready", None))
        state["a"] = [sx, sy, sz, sh]
        state["f"] = fl
        state["keys"] = keys
        mark("sync")
        if after_sync is not None:
            after_sync()
        if world == 1:
            _check(_fn("tree_view", bits, torch.float64)(dom._h, tree_ptrs), "tree_view")
            halos_single_rank(None, dom._info()[3], tree_ptrs)
        mark("find_halos_8_virtual_ranks")
        nw = dom.n_particles
        if state["nb"] is None or state["nb"].shape[0] < nw:
            cap = int(nw * 1.05) + 1024
            state["nb"] = torch.empty((cap, args.ngmax), dtype=torch.int32, device="cuda")
            state["nc"] = torch.empty(cap, dtype=torch.int32, device="cuda")
        dom.find_neighbors(sx, sy, sz, sh, args.ngmax, state["nb"], state["nc"])
        mark("find_neighbors")
        return nw

    # ---- warm-up: the first sync distributes the unsorted cloud and converges the trees (cold), the rest are steady
    sampler = ClockSampler(local_rank)  # before the cold step: 8 concurrent nvidia-smi need seconds to come up
    t0 = time.perf_counter()
    run_step()
    barrier()
    first_sync_s = time.perf_counter() - t0
    for _ in range(max(args.warmup, 3) - 1):
        run_step()
    barrier()

    def check_run(samples=64):
        """correctness evidence inside the run (after a steady warm-up step): particle conservation, SFC order, the
        fields still belong to their particles, and the neighbour counts of `samples` targets per rank against an
        all-pairs count over the assigned particles of ALL ranks"""
        sx, sy, sz, sh = state["a"]
        s_, e_ = dom.start_index, dom.end_index
        nw = e_ - s_
        k = state["keys"]
        out = dict(sorted_keys=bool((k[1:] >= k[:-1]).all().item()) if k.numel() > 1 else True)
        tot = torch.tensor([float(nw)], dtype=torch.float64, device="cuda")
        if args.fields:
            f0 = state["f"][0][s_:e_]
            out["fields_follow_particles"] = bool(all(torch.equal(state["f"][q][s_:e_], f0 * float(q + 1))
                                                      for q in range(1, args.fields)))
            tot = torch.cat([tot, f0.sum().reshape(1)])
        idx = torch.arange(samples, device="cuda", dtype=torch.int64) * (nw - 1) // max(samples - 1, 1) + s_
        tg = torch.stack([sx[idx], sy[idx], sz[idx], sh[idx]], 1)  # [samples, 4]
        mine = state["nc"][idx - s_].to(torch.float64)
        if world > 1:
            dist.all_reduce(tot)
            allt = [torch.empty_like(tg) for _ in range(world)]
            dist.all_gather(allt, tg)
            tg_all = torch.cat(allt, 0)
        else:
            tg_all = tg
        ax, ay, az = sx[s_:e_], sy[s_:e_], sz[s_:e_]
        cnt = torch.zeros(tg_all.shape[0], dtype=torch.float64, device="cuda")
        for q in range(tg_all.shape[0]):
            tx, ty, tz, th = (tg_all[q, c] for c in range(4))
            d2 = (ax - tx) ** 2 + (ay - ty) ** 2 + (az - tz) ** 2  # the reference's expression: plain squares, x + y + z
            cnt[q] = (d2 < 4.0 * th * th).sum()
        if world > 1:
            dist.all_reduce(cnt)
        expect = cnt[rank * samples:(rank + 1) * samples] - 1.0  # the target itself
        # u64 keys: exact.  u32 keys: the reference's coordinate -> cell conversion, floor(x * m) - xmin * m truncated
        # [ref: sfc/sfc.hpp sfc3D], can put a particle one cell off when the box origin is not a multiple of the cell
        # size; with 10 bits per dimension that is 1e-3 of the box, so its own tree search (which this library
        # reproduces bit for bit, see parity_in_run) misses a neighbour now and then: report the agreement instead
        dev_ = (expect - mine).abs()
        stats = torch.stack([(dev_ == 0).sum().to(torch.float64), dev_.max()])
        if world > 1:
            agree = stats[0:1].clone()
            worst = stats[1:2].clone()
            dist.all_reduce(agree)
            dist.all_reduce(worst, op=dist.ReduceOp.MAX)
            stats = torch.cat([agree, worst])
        n_equal, worst_dev = int(stats[0].item()), int(stats[1].item())
        if bits == 64:
            ok_nb = torch.tensor([float(n_equal == samples * world)], dtype=torch.float64, device="cuda")
        else:
            ok_nb = torch.tensor([float(n_equal >= 0.8 * samples * world and worst_dev <= 4)], dtype=torch.float64,
                                 device="cuda")
            out["neighbour_counts_equal"] = "%d of %d, largest deviation %d (u32 keys: see bench.py)" % (
                n_equal, samples * world, worst_dev)
        out["particles_conserved"] = bool(int(tot[0].item()) == world * n)
        if args.fields:
            ng = world * n
            out["id_sum_conserved"] = bool(tot[1].item() == ng * (ng - 1) / 2.0)
        out["neighbour_counts_vs_all_pairs"] = bool(ok_nb.item() == 1.0)
        out["sampled_targets"] = samples * world
        out["ok"] = all(v for k_, v in out.items() if isinstance(v, bool))
        return out

    check = None if args.no_check else check_run()
    barrier()

    # ---- device-resident timing: EXACTLY K steps between two events
    launches0 = pkg.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_acc, engine_acc = {}, {}
    barrier()
    wall0 = time.time()
    ev0.record()
    all_marks = []
    n_assigned = 0
    for _ in range(args.steps):
        marks = []
        n_assigned = run_step(marks)
        all_marks.append(marks)
        for k, v in dom.stage_ms().items():
            engine_acc[k] = engine_acc.get(k, 0.0) + v / args.steps
    ev1.record()
    barrier()
    wall1 = time.time()
    ms_total = ev0.elapsed_time(ev1)
    launches = pkg.launch_count() - launches0
    clocks = sampler.stop(wall0, wall1)
    for marks in all_marks:
        for (_, a), (name, b) in zip(marks[:-1], marks[1:]):
            stage_acc[name] = stage_acc.get(name, 0.0) + a.elapsed_time(b) / args.steps
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([float(n_assigned), float(dom.n_particles_with_halos - n_assigned)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms_step = float(t.item()) / args.steps
    n_global = int(cnt[0].item())
    value = n_global / (ms_step * 1e-3) / 1e6
    nc = state["nc"][:n_assigned]
    nc_mean = float(nc.to(torch.float64).mean().item())
    nc_stored = float(torch.clamp(nc, max=args.ngmax).to(torch.float64).mean().item())
    info = dom._info()

    # ---- end to end: pinned host -> device -> sync + findNeighbors -> pinned host, every step
    e2e = None
    if not args.no_e2e:
        nt = dom.n_particles_with_halos
        hin = [a.cpu().pin_memory() for a in state["a"]]
        hout = dict(keys=torch.empty(int(nt * 1.2) + 1024, dtype=torch.int64 if bits == 64 else torch.int32).pin_memory(),
                    arr=[torch.empty(int(nt * 1.2) + 1024, dtype=torch.float64).pin_memory() for _ in range(4)],
                    nc=torch.empty(int(n_assigned * 1.2) + 1024, dtype=torch.int32).pin_memory())
        moved = {"h2d": 0, "d2h": 0}

        d2h_stream, h2d_stream = torch.cuda.Stream(), torch.cuda.Stream()

        def upload(after=None):
            """x, y, z, h of the NEXT step from the pinned host buffers (on the upload stream; `after[k]` = event behind
            which host buffer k holds its data).  x, y, z first: box and keys need all of them; the engine waits for
            `h_ready` only where it first reads h, so box, keys and sort overlap the rest of the upload."""
            with torch.cuda.stream(h2d_stream):
                up = []
                for k in range(3):
                    if after is not None:
                        h2d_stream.wait_event(after[k])
                    up.append(hin[k].cuda(non_blocking=True))
                xyz_ready = torch.cuda.Event()
                xyz_ready.record()
                if after is not None:
                    h2d_stream.wait_event(after[3])
                up.append(hin[3].cuda(non_blocking=True))
                h_ready = torch.cuda.Event()
                h_ready.record()
            return dict(t=up, xyz_ready=xyz_ready, h_ready=h_ready, bytes=sum(t_.numel() * t_.element_size() for t_ in hin))

        def e2e_step():
            # the inputs of this step travel since the previous step's download delivered them to the host (double
            # buffering: PCIe up and down run next to findNeighbors); the first step uploads in its own time
            nxt = state.pop("next_up", None) or upload()
            torch.cuda.current_stream().wait_event(nxt["xyz_ready"])
            for t_ in nxt["t"]:
                t_.record_stream(torch.cuda.current_stream())
            state["a"] = nxt["t"]
            state["h_ready"] = nxt["h_ready"]
            moved["h2d"] = nxt["bytes"]

            def download_synced():
                # keys, x, y, z, h are final once sync returns: their download runs on a second stream while
                # findNeighbors computes (what a host-side client of Domain would do), and every array goes back up
                # for the next step as soon as it has arrived (the next step's input IS this step's output)
                m_ = dom.n_particles_with_halos
                ev = torch.cuda.Event()
                ev.record()
                d2h_stream.wait_event(ev)
                arrived = []
                with torch.cuda.stream(d2h_stream):
                    for k, (o, s_) in enumerate(zip(hout["arr"], state["a"])):
                        o[:m_].copy_(s_, non_blocking=True)
                        e_ = torch.cuda.Event()
                        e_.record()
                        arrived.append(e_)
                        hin[k] = o[:m_]
                    hout["keys"][:m_].copy_(state["keys"], non_blocking=True)
                state["next_up"] = upload(after=arrived)

            nw = run_step(after_sync=download_synced)
            m = dom.n_particles_with_halos
            hout["nc"][:nw].copy_(state["nc"][:nw], non_blocking=True)
            moved["d2h"] = m * (bits // 8) + 4 * m * 8 + nw * 4
            torch.cuda.synchronize()  # the host holds this step's results (and the next inputs are on the device)

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
        # what reached the host after the last step: sorted keys, the device's neighbour counts, the device's arrays
        m_l, nw_l = dom.n_particles_with_halos, dom.n_particles
        hk = hout["keys"][:m_l]
        e2e_ok = bool((hk[1:] >= hk[:-1]).all().item()) and bool(torch.equal(hout["nc"][:nw_l], state["nc"][:nw_l].cpu())) and \
            bool(torch.equal(hout["arr"][0][:m_l], state["a"][0].cpu()))
        e2e = dict(value=n_global / (ms_e2e * 1e-3) / 1e6, unit=UNIT, h2d_bytes_per_step=int(moved["h2d"]),
                   d2h_bytes_per_step=int(moved["d2h"]), ms_per_step=ms_e2e,
                   outputs="keys, x, y, z, h (assigned + halos, SFC order) and neighbour counts; the neighbour lists "
                           "(%d x %d x 4 B per rank) stay device-resident as for Domain<...,GpuTag> clients" %
                           (n_assigned, args.ngmax),
                   host_results_ok=e2e_ok,
                   pipeline="double-buffered: the download of a step's arrays and their upload as the next step's "
                            "inputs run on two copy streams next to findNeighbors; every step waits for its own "
                            "inputs and ends with its neighbour counts on the host")

    if rank != 0:
        dom.close()
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel + per-stage tables
    peak, peak_src = load_peaks()
    sK, sT = bits // 8, 8
    npart = n_assigned
    alg = {
        "sfc_keys": npart * (3 * sT + 2 * sK),
        "sort": npart * (sK + sK * 2 * (sK + 4) - 4),     # histogram read + sK passes r/w of (key, u32); pass 0 reads no value
        "find_neighbors": npart * (4 * sT + 4) + 4.0 * nc_stored * npart,
    }
    stages = {}
    for name, ms in list(stage_acc.items()) + [("sync." + k, v) for k, v in engine_acc.items()]:
        d = dict(ms=round(ms, 4), share=round(ms / ms_step, 4))
        key = name.replace("sync.", "")
        if key in alg:
            gbs = alg[key] / (ms * 1e-3) / 1e9
            d.update(alg_bytes=int(alg[key]), gbs=round(gbs, 1), frac=round(gbs / peak, 4))
        stages[name] = d
    ach = alg["find_neighbors"] / (stage_acc["find_neighbors"] * 1e-3) / 1e9
    roofline = dict(bound="hbm", kernel="findNeighborsKernel<double>", achieved=round(ach, 1), peak=peak, unit="GB/s",
                    frac=round(ach / peak, 4), traffic=NEIGHBOR_TRAFFIC_BYTES, peak_source=peak_src,
                    alg_bytes_per_launch=int(alg["find_neighbors"]), launch_ms=round(stage_acc["find_neighbors"], 4),
                    note="issue-bound, not HBM-bound: see DESIGN.md section 4; traffic is the ncu dram read+write of one "
                         "launch at the N=1 default size (profiles/), null otherwise")
    if not (world == 1 and n == 32_000_000 and args.dist == "uniform"):
        roofline["traffic"] = None

    cpu = None
    if not args.no_cpu and world == 1:
        # bounded sample of the same workload through the reference's own Domain (the full-size run is --impl reference)
        saved = args.n
        args.n = min(args.n, args.cpu_n)
        cpu, _, rd = time_reference(args, 1, 2, 1, args.cpu_n, full_n=saved)
        rd.close()
        args.n = saved
        if not args.no_check:
            check = dict(check or {}, parity_in_run=parity_in_run(args))
            check["ok"] = bool(check.get("ok", True) and check["parity_in_run"]["ok"])

    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                ms_per_step=ms_step, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64",
                data="synthetic", impl="b200",
                config=config,
                observed=dict(particles_global=n_global, halos_global=int(cnt[1].item()), leaves_rank0=int(info[3]),
                              global_leaves=int(info[5]), neighbours_mean=round(nc_mean, 2),
                              particles_per_leaf=round(n_assigned / max(int(info[3]), 1), 1),
                              first_sync_s=round(first_sync_s, 3),
                              halo_exchange=("n/a (1 rank)" if world == 1 else
                                             "pack-and-push over CUDA IPC peer memory" if args.halo_push else
                                             "gather per array + grouped ncclSend/ncclRecv"),
                              halo_discovery=("n/a (1 rank)" if world == 1 else
                                              "reference procedure on the multi-rank focus tree (halo mode 1)" if args.halo_mode
                                              else "fast path: surface boxes against assigned leaves (halo mode 0)"),
                              parallelism="1 rank" if world == 1 else
                              "%d ranks, SFC-range sharding, NCCL all-reduce of global leaf counts, grouped NCCL "
                              "send/recv for particle and halo exchange" % world),
                check=check, clocks=clocks, e2e=e2e, gpu_launches=int(launches), roofline=roofline, stages=stages,
                cpu_baseline=cpu)
    print(json.dumps(line))
    dom.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
