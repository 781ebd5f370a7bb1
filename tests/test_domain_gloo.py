"""world_size-2 check (gloo, CPU) of the host side of the sharded path: two processes hold different particle slabs,
build the replicated global tree the way the engine does (local counts -> all-reduce SUM), call the library's
host assignment, derive send ranges, exchange, and end with the same decomposition a single process computes on the
union [ref: domain/domain.hpp:262-300 distribute, domain/domaindecomp.hpp:176-192 createSendRanges,
tree/update_mpi.hpp:50-80 updateOctreeGlobal].  The device kernels are not involved; the oracle stands in for them."""
import ctypes as C
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_per_rank, bucket, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    from oracle import loader
    import cornerstone_octree_b200 as pkg
    import gen

    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    orc = loader.load("port")
    lib = pkg._lib.lib

    def p(a):
        return a.ctypes.data_as(C.c_void_p) if a is not None else None

    old_bounds = None
    for step in range(2):
        # every rank owns an unrelated slab, like bench.py's per-rank clouds; second step: drifted distribution
        keys = np.sort(gen.gaussian_keys(n_per_rank + 1000 * rank, np.uint64, seed=10 * step + rank))
        # replicated global tree: converge leaf counts through the all-reduce
        leaves = np.array([0, 1 << 63], dtype=np.uint64)
        counts = np.zeros(1, dtype=np.uint32)
        for _ in range(30):
            local = orc.compute_node_counts(leaves, keys)
            t = torch.from_numpy(local.astype(np.int64))
            dist.all_reduce(t)
            counts = np.minimum(t.numpy(), 0xFFFFFFFF).astype(np.uint32)
            ops = orc.rebalance_decision(leaves, counts, bucket)
            if ops[1]:
                break
            leaves = orc.rebalance_tree(leaves, ops[0])[0]
        nb = np.zeros(world + 1, dtype=np.uint64)
        rc = np.zeros(world, dtype=np.uint32)
        st = lib.cstone_b200_host_assignment_k64(world, p(leaves), p(counts), len(counts), p(old_bounds), p(nb), p(rc))
        assert st == 0
        old_bounds = nb.copy()
        # send ranges by lower bound, exchange with point-to-point messages
        cuts = np.searchsorted(keys, nb, side="left")
        cuts[-1] = keys.size
        mine = [keys[cuts[rank]:cuts[rank + 1]]]
        reqs = []
        for peer in range(world):
            if peer != rank:
                out = torch.from_numpy(keys[cuts[peer]:cuts[peer + 1]].astype(np.int64))
                size = torch.tensor([out.numel()])
                reqs.append(dist.isend(size, peer))
                reqs.append(dist.isend(out, peer))
        for peer in range(world):
            if peer != rank:
                size = torch.zeros(1, dtype=torch.int64)
                dist.recv(size, peer)
                buf = torch.zeros(int(size), dtype=torch.int64)
                dist.recv(buf, peer)
                mine.append(buf.numpy().astype(np.uint64))
        for r in reqs:
            r.wait()
        assigned = np.sort(np.concatenate(mine))
        assert assigned.size == int(rc[rank])
        assert assigned.size == 0 or (assigned[0] >= nb[rank] and assigned[-1] < nb[rank + 1])
        np.savez(os.path.join(out_dir, "r%d_s%d.npz" % (rank, step)), leaves=leaves, counts=counts, bounds=nb,
                 rank_counts=rc, assigned=assigned, keys=keys)
    dist.destroy_process_group()


def _compute_octree_global(orc, keys, bucket):
    leaves = np.array([0, 1 << 63], dtype=np.uint64)
    for _ in range(30):
        counts = orc.compute_node_counts(leaves, keys)
        ops = orc.rebalance_decision(leaves, counts, bucket)
        if ops[1]:
            break
        leaves = orc.rebalance_tree(leaves, ops[0])[0]
    return leaves, counts


def test_two_rank_decomposition(tmp_path, port):
    import torch.multiprocessing as mp
    world, n, bucket = 2, 30000, 512
    mp.spawn(_worker, args=(world, _free_port(), n, bucket, str(tmp_path)), nprocs=world, join=True)
    import test_domain_host as th
    old = None
    for step in range(2):
        r = [np.load(os.path.join(tmp_path, "r%d_s%d.npz" % (q, step))) for q in range(world)]
        # both ranks hold the same replicated tree and assignment
        for k in ("leaves", "counts", "bounds", "rank_counts"):
            assert np.array_equal(r[0][k], r[1][k])
        union = np.sort(np.concatenate([r[q]["keys"] for q in range(world)]))
        # ... which is the single-process tree of the union with the single-process assignment
        leaves, counts = _compute_octree_global(port, union, bucket)
        assert np.array_equal(leaves, r[0]["leaves"]) and np.array_equal(counts, r[0]["counts"])
        b, rc = th.assignment(world, leaves, counts, old=old)
        assert np.array_equal(b, r[0]["bounds"]) and np.array_equal(rc, r[0]["rank_counts"])
        old = b
        # conservation: the assigned sets partition the union in SFC order
        assert np.array_equal(np.concatenate([r[q]["assigned"] for q in range(world)]), union)
