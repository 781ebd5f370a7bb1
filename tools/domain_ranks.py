"""Multi-rank check of the Domain engine, one process per GPU:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/domain_ranks.py
Every rank starts with a slab of an unsorted global cloud.  After sync: (1) no particle is lost or duplicated,
(2) every rank's assigned keys are sorted and inside its SFC range, ranks hold nearly equal counts, (3) halos are
sufficient: the neighbour lists of the assigned particles, translated to global ids, equal the lists of a single
global search by the CPU oracle, (4) exchange_halos delivers the owner's values, (5) a second sync after a small
displacement keeps all of this."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200.domain import Domain
    from oracle import loader
    import gen
    import pipeline
    n_glob = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    cloud = sys.argv[2] if len(sys.argv) > 2 else "gaussian"
    ngmax = 200
    if cloud == "gaussian":
        x, y, z = gen.gaussian(n_glob, (-1.0, 1.0), np.float64)
    else:
        x, y, z = gen.uniform(n_glob, (-1.0, 1.0), np.float64)
    rng = np.random.default_rng(3)
    h = 0.5 * 2.0 * np.cbrt(3 * 50 / (4 * np.pi * n_glob)) * (0.8 + 0.4 * rng.random(n_glob))
    gid = np.arange(n_glob, dtype=np.int64)
    lo, hi = rank * n_glob // world, (rank + 1) * n_glob // world
    dom = Domain(rank, world, 512, 32, 0.5, ops.make_box((-1.0, 1.0)))
    dom.init_comm_from_torch()
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    ok = True

    def check(tag, keys, sx, sy, sz, sh, sgid, X, Y, Z, H):
        nonlocal ok
        s, e, nt = dom.start_index, dom.end_index, dom.n_particles_with_halos
        k = keys.cpu().numpy().view(np.uint64)
        g = sgid.cpu().numpy()
        # (1) conservation
        counts = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
        dist.all_gather(counts, torch.tensor([e - s], dtype=torch.int64, device="cuda"))
        counts = [int(c.item()) for c in counts]
        allg = [torch.empty(c, dtype=torch.int64, device="cuda") for c in counts]
        dist.all_gather(allg, sgid[s:e].contiguous())
        allg = torch.cat(allg).cpu().numpy()
        c1 = np.array_equal(np.sort(allg), np.arange(n_glob))
        # (2) sortedness, range, balance
        leaves, gcounts, bounds = dom.global_tree()
        c2 = bool(np.all(k[1:] >= k[:-1])) and (e == s or (k[s] >= bounds[rank] and k[e - 1] < bounds[rank + 1]))
        c2 = c2 and max(counts) - min(counts) <= 2 * 512 + 2
        # coordinates really belong to the ids (assigned and halos)
        dist.barrier()
        dom.exchange_halos([sgid])
        g = sgid.cpu().numpy()
        c4 = np.array_equal(sx.cpu().numpy(), X[g]) and np.array_equal(sh.cpu().numpy(), H[g])
        # (3) neighbours vs one global search
        nb, nc = dom.find_neighbors(sx, sy, sz, sh, ngmax)
        nb, nc = nb.cpu().numpy(), nc.cpu().numpy()
        port = loader.load("port")
        b = dom.box()
        obox = loader.make_box([b.limits[i] for i in range(6)])
        r = pipeline.run(port, 1, X, Y, Z, H, obox, np.uint64, 32, ngmax)
        inv = np.empty(n_glob, dtype=np.int64)
        inv[r["order"]] = np.arange(n_glob)          # global id -> position in the oracle's sorted arrays
        c3 = True
        step = max(1, (e - s) // 4000)
        for i in range(s, e, step):
            mine = sorted(g[nb[i - s, :min(nc[i - s], ngmax)]].tolist())
            o = inv[g[i]]
            ref = sorted(r["order"][r["neighbors"][o, :min(r["neighbors_count"][o], ngmax)]].tolist())
            if nc[i - s] != r["neighbors_count"][o] or (nc[i - s] <= ngmax and mine != ref):
                c3 = False
                print("rank", rank, tag, "neighbour mismatch at", i, nc[i - s], r["neighbors_count"][o], flush=True)
                break
        print("rank %d %s: assigned %d halos %d+%d | conservation %s sorted/range/balance %s halos-sufficient %s "
              "exchange %s" % (rank, tag, e - s, s, nt - e, c1, c2, c3, c4), flush=True)
        ok = ok and c1 and c2 and c3 and c4

    keys, sx, sy, sz, sh, (sgid,) = dom.sync(dev(x[lo:hi]), dev(y[lo:hi]), dev(z[lo:hi]), dev(h[lo:hi]), [dev(gid[lo:hi])])
    check("sync1", keys, sx, sy, sz, sh, sgid, x, y, z, h)
    # move everything a little (same displacement field on every rank), sync again
    disp = 0.02 * np.random.default_rng(11).standard_normal((3, n_glob))
    x2, y2, z2 = np.clip(x + disp[0], -1, 1), np.clip(y + disp[1], -1, 1), np.clip(z + disp[2], -1, 1)
    g = sgid.cpu().numpy()
    keys, sx, sy, sz, sh, (sgid,) = dom.sync(dev(x2[g]), dev(y2[g]), dev(z2[g]), sh, [sgid])
    check("sync2", keys, sx, sy, sz, sh, sgid, x2, y2, z2, h)
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("DOMAIN_RANKS_OK" if int(t.item()) == 1 else "DOMAIN_RANKS_FAILED", flush=True)
    dom.close()
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
