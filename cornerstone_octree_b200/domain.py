"""Host-side mirror of cstone::Domain<KeyType,T,GpuTag> [ref: domain/domain.hpp:66-699] on torch CUDA tensors.

Every method forwards to the domain engine of libcstone_b200.so (include/cstone_b200_domain.h); torch provides the
device memory of the caller-owned particle arrays and, for multi-GPU runs, the bootstrap of the NCCL communicator
(the unique id is broadcast through torch.distributed, all data-path communication is NCCL inside the library)."""
import ctypes as C

import numpy as np
import torch

from . import _lib, ops
from ._lib import BoxF32, BoxF64, CstoneError, lib


def _fn(name, bits, fdtype):
    ts = "f32" if fdtype == torch.float32 else "f64"
    return getattr(lib, "cstone_b200_domain_%s_k%d_%s" % (name, bits, ts))


def _check(st, where):
    if st != 0:
        raise CstoneError(int(st), "cstone_b200_domain_" + where)


def _row_bytes(t):
    """bytes per particle of a 1-D or row-per-particle tensor"""
    n = 1
    for d in t.shape[1:]:
        n *= int(d)
    return t.element_size() * n


class Domain:
    """Domain(rank, num_ranks, bucket_size, bucket_size_focus, theta, box)  — same arguments as the reference"""

    def __init__(self, rank, num_ranks, bucket_size, bucket_size_focus, theta=0.5, box=None, key_bits=64,
                 dtype=torch.float64, sfc_kind=ops.HILBERT):
        if bucket_size < bucket_size_focus:
            # [ref: domain/domain.hpp:108-112]
            raise RuntimeError("The bucket size of the global tree must not be smaller than the bucket size of the "
                               "focused tree")
        self.rank, self.num_ranks, self.bits, self.fd = rank, num_ranks, key_bits, dtype
        box = box if box is not None else ops.make_box((0.0, 1.0), dtype=dtype)
        self._h = C.c_void_p(_fn("create", key_bits, dtype)(rank, num_ranks, bucket_size, bucket_size_focus, theta,
                                                          C.byref(box)))
        if not self._h:
            raise RuntimeError("cstone_b200_domain_create failed")
        self._comm = None
        self.sfc_kind = sfc_kind
        if sfc_kind != ops.HILBERT:  # the reference's Domain hard-wires Hilbert; Morton is an extension
            _check(_fn("set_sfc_kind", key_bits, dtype)(self._h, int(sfc_kind)), "set_sfc_kind")

    def close(self):
        if getattr(self, "_h", None):
            _fn("destroy", self.bits, self.fd)(self._h)
            self._h = None
        if self._comm:
            lib.cstone_b200_nccl_comm_destroy(self._comm)
            self._comm = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- multi-GPU bootstrap
    def init_comm_from_torch(self):
        """create the NCCL communicator of this Domain; the 128-byte unique id travels over torch.distributed"""
        import torch.distributed as dist
        if self.num_ranks == 1:
            return
        ident = torch.zeros(128, dtype=torch.uint8)
        if self.rank == 0:
            buf = (C.c_ubyte * 128)()
            _check(lib.cstone_b200_nccl_unique_id(buf), "nccl_unique_id")
            ident = torch.tensor(list(buf), dtype=torch.uint8)
        dev = ident.cuda() if dist.get_backend() == "nccl" else ident
        dist.broadcast(dev, src=0)
        raw = bytes(dev.cpu().tolist())
        comm = C.c_void_p()
        _check(lib.cstone_b200_nccl_comm_init(raw, self.num_ranks, self.rank, C.byref(comm)), "nccl_comm_init")
        self._comm = comm
        _check(_fn("set_comm", self.bits, self.fd)(self._h, comm), "set_comm")

    def set_local_world(self, world):
        """in-process transport: this Domain is one of `world.num_ranks` thread-ranks sharing the GPU (see LocalWorld)"""
        self._world = world  # keep it alive
        _check(_fn("set_local_world", self.bits, self.fd)(self._h, world.handle), "set_local_world")

    # ---- sync
    def sync_grav(self, x, y, z, h, m, props=()):
        """[ref: domain/domain.hpp:245-325 syncGrav] like sync, with the focus tree refined by the vector MAC around the
        expansion centres; m (masses, same dtype as x) travels like a property and is returned after h:
        (keys, x, y, z, h, m, props).  Needs set_halo_mode(1) and more than one rank."""
        out = self.sync(x, y, z, h, (m,) + tuple(props), _entry="sync_grav_begin")
        return out[0], out[1], out[2], out[3], out[4], out[5][0], out[5][1:]

    def let_centers(self):
        """after sync_grav: expansion centres [numNodes, 4] = x, y, z, MAC radius squared, and the MAC marks [numNodes]
        of the focus tree (torch tensors that alias engine memory, valid until the next sync)"""
        fd = C.c_double if self.fd == torch.float64 else C.c_float
        c, m, n = C.POINTER(fd)(), C.c_char_p(), C.c_int32(0)
        cp, mp = C.c_void_p(0), C.c_void_p(0)
        _check(_fn("let_centers", self.bits, self.fd)(self._h, C.byref(cp), C.byref(mp), C.byref(n)), "let_centers")
        return _alias(cp.value, 4 * n.value, self.fd).view(n.value, 4), _alias(mp.value, n.value, torch.int8)

    # ---- globalFocusExchange for client node quantities [ref: focus/octree_focus_mpi.hpp:276-366, :762-784]
    def focus_peer_exchange(self, node_q):
        """cells of the focus tree in a peer's SFC range take the owner's values (node_q: one row per tree node)"""
        _check(_fn("focus_peer_exchange", self.bits, self.fd)(self._h, ops._p(node_q), _row_bytes(node_q)), "focus_peer_exchange")

    def focus_global_gather(self, node_q):
        """one row per leaf of the replicated global tree: this rank's leaves from node_q, the others from their owners"""
        n_global = self._info()[5]
        out = torch.zeros((n_global,) + tuple(node_q.shape[1:]), dtype=node_q.dtype, device=node_q.device)
        _check(_fn("focus_global_gather", self.bits, self.fd)(self._h, ops._p(node_q), _row_bytes(node_q), ops._p(out)),
               "focus_global_gather")
        return out

    def focus_global_extract(self, global_node_q, node_q):
        """cells of the focus tree that belong to no peer read their rows from global_node_q (one row per global tree node)"""
        _check(_fn("focus_global_extract", self.bits, self.fd)(self._h, ops._p(global_node_q), _row_bytes(node_q), ops._p(node_q)),
               "focus_global_extract")

    def global_octree(self):
        """the replicated global tree linked on the device (after focus_global_gather or sync_grav): dict of tensors +
        the level ranges as a host list"""
        nl, nn = C.c_int32(0), C.c_int32(0)
        ptrs = (C.c_void_p * 6)()
        ml = ops.max_level(self.bits)
        lr = (C.c_int32 * (ml + 2))()
        _check(_fn("global_octree", self.bits, self.fd)(self._h, C.byref(nl), C.byref(nn), ptrs, lr), "global_octree")
        L, M = nl.value, nn.value
        kt, i32 = ops.KEY_TORCH[self.bits], torch.int32
        shapes = [("prefixes", kt, M), ("child_offsets", i32, M + 1), ("parents", i32, max(1, (M - 1) // 8)),
                  ("level_range", i32, ml + 2), ("internal_to_leaf", i32, M), ("leaf_to_internal", i32, M)]
        out = {name: _alias(p, n, dt) for (name, dt, n), p in zip(shapes, ptrs)}
        out.update(num_leaves=L, num_nodes=M, level_range_host=[int(v) for v in lr])
        return out

    def sync(self, x, y, z, h, props=(), fields_ready=None, _entry="sync_begin"):
        """[ref: domain/domain.hpp:196-243] returns (keys, x, y, z, h, props) as NEW tensors of the new local size:
        assigned particles in [start_index, end_index), halos elsewhere, everything in SFC order.
        fields_ready: optional torch.cuda.Event after which h and props are valid (e.g. the end of their upload on
        another stream); x, y, z must be valid at the call, the engine waits for the event before it first reads
        h / props, so that box, keys and sort overlap the rest of the upload."""
        for t in (x, y, z, h) + tuple(props):
            if not t.is_cuda or not t.is_contiguous():
                raise ValueError("Domain.sync operates on contiguous CUDA tensors (no CPU fallback)")
        n_in = x.numel()
        k = len(props)
        parr = (C.c_void_p * max(k, 1))(*[p.data_ptr() for p in props])
        # a property is one value per particle or one ROW per particle (2-D tensor): 4, 8, 12 or 16 bytes per particle
        # [ref: domain/buffer_description.hpp:98-110: util::array elements travel like scalars]
        pbytes = (C.c_int * max(k, 1))(*[_row_bytes(p) for p in props])
        if fields_ready is not None:
            _check(_fn("set_fields_ready_event", self.bits, self.fd)(self._h, C.c_void_p(fields_ready.cuda_event)),
                   "set_fields_ready_event")
        n_new = _fn(_entry, self.bits, self.fd)(self._h, n_in, ops._p(x), ops._p(y), ops._p(z), ops._p(h), k, parr, pbytes)
        if n_new < 0:
            raise CstoneError(int(n_new), "cstone_b200_domain_" + _entry)
        dev = x.device
        keys = torch.empty(n_new, dtype=ops.KEY_TORCH[self.bits], device=dev)
        out = [torch.empty(n_new, dtype=x.dtype, device=dev) for _ in range(4)]
        pout = [torch.empty((n_new,) + tuple(p.shape[1:]), dtype=p.dtype, device=dev) for p in props]
        oarr = (C.c_void_p * max(k, 1))(*[p.data_ptr() for p in pout])
        _check(_fn("sync_finish", self.bits, self.fd)(self._h, ops._p(keys), ops._p(out[0]), ops._p(out[1]), ops._p(out[2]),
                                                    ops._p(out[3]), oarr), "sync_finish")
        return keys, out[0], out[1], out[2], out[3], pout

    def exchange_halos(self, arrays):
        """[ref: domain/domain.hpp:381-386] fill the halo ranges of `arrays` (length n_particles_with_halos) in place"""
        k = len(arrays)
        if k == 0:
            return
        arr = (C.c_void_p * k)(*[a.data_ptr() for a in arrays])
        eb = (C.c_int * k)(*[_row_bytes(a) for a in arrays])
        _check(_fn("exchange_halos", self.bits, self.fd)(self._h, k, arr, eb), "exchange_halos")

    def reapply_sync(self, arrays):
        """[ref: domain/domain.hpp:334-378 reapplySync] arrays in the layout BEFORE the last sync -> new tensors of the
        new local size with the assigned range filled (halos: exchange_halos)"""
        k = len(arrays)
        if k == 0:
            return []
        n_new = self.n_particles_with_halos
        out = [torch.zeros((n_new,) + tuple(a.shape[1:]), dtype=a.dtype, device=a.device) for a in arrays]
        iarr = (C.c_void_p * k)(*[a.data_ptr() for a in arrays])
        oarr = (C.c_void_p * k)(*[a.data_ptr() for a in out])
        eb = (C.c_int * k)(*[_row_bytes(a) for a in arrays])
        _check(_fn("reapply_sync", self.bits, self.fd)(self._h, k, iarr, eb, oarr), "reapply_sync")
        return out

    # ---- accessors [ref: domain/domain.hpp:389-437]
    def _info(self):
        info = (C.c_int64 * 8)()
        _check(_fn("info", self.bits, self.fd)(self._h, info), "info")
        return list(info)

    @property
    def start_index(self):
        return self._info()[0]

    @property
    def end_index(self):
        return self._info()[1]

    def set_end_index(self, i):
        """[ref: domain/domain.hpp:396 setEndIndex] particles written behind the assigned range join it at the next sync"""
        _check(_fn("set_end_index", self.bits, self.fd)(self._h, int(i)), "set_end_index")

    def set_halo_factor(self, factor):
        """[ref: domain/domain.hpp:412 setHaloFactor] halo search radius = 2 * factor * max h of a leaf"""
        _check(_fn("set_halo_factor", self.bits, self.fd)(self._h, float(factor)), "set_halo_factor")

    def cells(self):
        """[ref: domain/domain.hpp:404-406 startCell / endCell] leaves of the search tree inside the assigned SFC range"""
        c = (C.c_int32 * 2)()
        _check(_fn("cells", self.bits, self.fd)(self._h, c), "cells")
        return int(c[0]), int(c[1])

    @property
    def n_particles(self):
        i = self._info()
        return i[1] - i[0]

    @property
    def n_particles_with_halos(self):
        return self._info()[2]

    STAGES = ["box", "sfc_keys", "sort", "global_tree", "exchange", "resort", "halo_discovery", "halo_exchange",
              "search_tree", "link_tree"]

    def stage_ms(self):
        """device time (ms) of the stages of the last sync, measured with CUDA events inside the engine"""
        ms = (C.c_float * 10)()
        _check(_fn("stage_ms", self.bits, self.fd)(self._h, ms), "stage_ms")
        return dict(zip(self.STAGES, [float(v) for v in ms]))

    def box(self):
        b = BoxF32() if self.fd == torch.float32 else BoxF64()
        _check(_fn("box", self.bits, self.fd)(self._h, C.byref(b)), "box")
        return b

    def set_halo_mode(self, mode):
        """0: fast halo discovery (default); 1: the reference's procedure on the multi-rank focus tree
        [ref: domain/domain.hpp:217-243] -- halo flags, layout, start/end index and the search tree are the reference's"""
        _check(_fn("set_halo_mode", self.bits, self.fd)(self._h, int(mode)), "set_halo_mode")

    def set_halo_push(self, on):
        """transport of the halo exchange: True (default) = pack-and-push over peer memory, False = gather + send/recv"""
        _check(_fn("set_halo_push", self.bits, self.fd)(self._h, 1 if on else 0), "set_halo_push")

    def let_view(self):
        """halo mode 1: host copies of this rank's focus tree after a sync: leaves, leaf counts, halo flags, layout,
        leaf index range of every rank (focus assignment) and the peer ranks"""
        kt = C.c_uint64 if self.bits == 64 else C.c_uint32
        n, npeers = C.c_int32(0), C.c_int32(0)
        l, c, f, lay = C.POINTER(kt)(), C.POINTER(C.c_uint32)(), C.POINTER(C.c_int32)(), C.POINTER(C.c_uint32)()
        asg = np.zeros(2 * self.num_ranks, dtype=np.int32)
        peers = np.zeros(self.num_ranks, dtype=np.int32)
        _check(_fn("let_view", self.bits, self.fd)(self._h, C.byref(n), C.byref(l), C.byref(c), C.byref(f), C.byref(lay),
                                                   asg.ctypes.data_as(C.c_void_p), peers.ctypes.data_as(C.c_void_p),
                                                   C.byref(npeers)), "let_view")
        L = n.value
        kd = np.uint64 if self.bits == 64 else np.uint32
        return dict(leaves=np.ctypeslib.as_array(l, (L + 1,)).astype(kd), counts=np.ctypeslib.as_array(c, (L,)).copy(),
                    flags=np.ctypeslib.as_array(f, (L,)).copy(), layout=np.ctypeslib.as_array(lay, (L + 1,)).copy(),
                    assignment=asg.reshape(-1, 2), peers=peers[:npeers.value].copy())

    def global_tree(self):
        """replicated global leaf array, its counts and the rank boundaries (numpy, host)"""
        kt = C.c_uint64 if self.bits == 64 else C.c_uint32
        l, c, a = C.POINTER(kt)(), C.POINTER(C.c_uint32)(), C.POINTER(kt)()
        _check(_fn("global_tree", self.bits, self.fd)(self._h, C.byref(l), C.byref(c), C.byref(a)), "global_tree")
        L = self._info()[5]
        kd = np.uint64 if self.bits == 64 else np.uint32
        return (np.ctypeslib.as_array(l, (L + 1,)).astype(kd), np.ctypeslib.as_array(c, (L,)).copy(),
                np.ctypeslib.as_array(a, (self.num_ranks + 1,)).astype(kd))

    def tree_arrays(self):
        """device views (torch tensors that alias engine memory, valid until the next sync) of the search tree"""
        ptrs = (C.c_void_p * 10)()
        _check(_fn("tree_view", self.bits, self.fd)(self._h, ptrs), "tree_view")
        i = self._info()
        L, M = i[3], i[4]
        ml = ops.max_level(self.bits)
        kt, i32 = ops.KEY_TORCH[self.bits], torch.int32
        shapes = [("leaves", kt, L + 1), ("counts", i32, L), ("layout", i32, L + 1), ("prefixes", kt, M),
                  ("child_offsets", i32, M + 1), ("internal_to_leaf", i32, M), ("leaf_to_internal", i32, M),
                  ("level_range", i32, ml + 2), ("centers", self.fd, 3 * M), ("sizes", self.fd, 3 * M)]
        out = {}
        for (name, dt, n), p in zip(shapes, ptrs):
            out[name] = _alias(p, n, dt)
        return out

    def find_neighbors(self, x, y, z, h, ngmax, neighbors=None, neighbors_count=None):
        """[ref: findneighbors.hpp:167-178] over [start_index, end_index) with the Domain's tree"""
        nw = self.n_particles
        if neighbors is None:
            neighbors = torch.empty((nw, ngmax), dtype=torch.int32, device=x.device)
        if neighbors_count is None:
            neighbors_count = torch.empty(nw, dtype=torch.int32, device=x.device)
        _check(_fn("find_neighbors", self.bits, self.fd)(self._h, ops._p(x), ops._p(y), ops._p(z), ops._p(h), ngmax,
                                                       ops._p(neighbors), ops._p(neighbors_count)), "find_neighbors")
        return neighbors, neighbors_count


def _alias(ptr, n, dtype):
    """torch tensor over `n` elements of device memory owned by the engine (copied: views must not outlive a sync)"""
    out = torch.empty(n, dtype=dtype, device="cuda")
    if n:
        nbytes = n * out.element_size()
        rt = C.CDLL("libcudart.so")
        rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        st = rt.cudaMemcpy(C.c_void_p(out.data_ptr()), C.c_void_p(ptr), nbytes, 3)
        if st != 0:
            raise RuntimeError("cudaMemcpy failed: %d" % st)
    return out


class LocalWorld:
    """P ranks as P host threads of this process on ONE GPU (include/cstone_b200_domain.h, in-process transport): a
    checking tool for the multi-rank path on single-GPU machines.  Usage:
        world = LocalWorld(P); doms = [Domain(r, P, ...) for r in range(P)]; [d.set_local_world(world) for d in doms]
        results = world.run(lambda r: doms[r].sync(...))      # one thread per rank, collectives meet inside"""

    def __init__(self, num_ranks):
        self.num_ranks = num_ranks
        self.handle = C.c_void_p(lib.cstone_b200_local_world_create(num_ranks))
        if not self.handle:
            raise RuntimeError("cstone_b200_local_world_create failed")

    def run(self, fn):
        """fn(rank) on one thread per rank (ctypes releases the GIL inside the library); returns the results in rank
        order and re-raises the first exception"""
        import threading
        out, err = [None] * self.num_ranks, [None] * self.num_ranks
        dev = torch.cuda.current_device()

        def body(r):
            try:
                torch.cuda.set_device(dev)
                out[r] = fn(r)
            except BaseException as e:  # noqa: BLE001  (reported to the caller below)
                err[r] = e

        threads = [threading.Thread(target=body, args=(r,)) for r in range(self.num_ranks)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        for e in err:
            if e is not None:
                raise e
        return out

    def close(self):
        if self.handle:
            lib.cstone_b200_local_world_destroy(self.handle)
            self.handle = None
