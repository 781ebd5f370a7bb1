"""Single-rank hot path with preallocated device buffers: the sequence of C-ABI calls one
`Domain<KeyType,T,GpuTag>::sync` + `findNeighbors` makes for the particles of one rank
[ref: domain/domain.hpp:196-243 sync, domain/assignment_gpu.cuh:92-118 (keys + sort), domain/layout.hpp:202-239
(gatherArrays), tree/update_gpu.cuh:58-82 (leaf array), tree/octree_gpu.cu:152-176 (linking),
halos/halos.hpp:127-189 (discover), findneighbors.hpp:167-188].

Nothing here computes: every stage is one call into libcstone_b200.so on torch's current stream; torch provides
device memory only.  Used by bench.py, __graft_entry__.smoke() and the full-size property tests."""
import ctypes as C

import numpy as np
import torch

from . import ops
from ._lib import HILBERT, call

STAGES = ["sfc_keys", "sort", "gather", "leaf_array", "link_tree", "geo_centers", "layout", "halo_radii", "find_halos",
          "find_neighbors"]


class HotPath:
    """owns every buffer of one rank's sync + findNeighbors pass over n particles (x,y,z,h of dtype `fdtype`)"""

    def __init__(self, n, key_bits=64, fdtype=torch.float64, bucket=64, ngmax=150, kind=HILBERT, virtual_ranks=8,
                 leaf_capacity=None, device="cuda", with_neighbors=True):
        self.n, self.bits, self.fd, self.bucket, self.ngmax, self.kind = n, key_bits, fdtype, bucket, ngmax, kind
        self.vranks = virtual_ranks
        self.dev = device
        kt = ops.KEY_TORCH[key_bits]
        self.keys = torch.zeros(n, dtype=kt, device=device)
        self.sorter = ops.SfcSorter(device)
        self.sorted = [torch.empty(n, dtype=fdtype, device=device) for _ in range(4)]
        cap = leaf_capacity or max(1024, int(n / max(bucket // 8, 1)))
        self.leaves = ops.LeafArray(key_bits, capacity=cap, device=device)
        self.ws = ops.Workspace(device)
        self.with_neighbors = with_neighbors
        self.neighbors = torch.empty((n, ngmax), dtype=torch.int32, device=device) if with_neighbors else None
        self.neighbors_count = torch.empty(n, dtype=torch.int32, device=device)
        self.octree = None
        self.centers = self.sizes = self.layout = self.radii = self.flags = None
        self.first_call = True
        self.events = None

    # stage markers ------------------------------------------------------------------------------
    def _mark(self, name):
        if self.events is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.events.append((name, e))

    def _alloc_tree_buffers(self, L):
        M = L + (L - 1) // 7
        if getattr(self, "_oct_store", None) is not None and self._oct_store["prefixes"].numel() >= M:
            return
        dev, i32 = self.dev, torch.int32
        cap = int(M * 1.05) + 64
        kt = ops.KEY_TORCH[self.bits]
        self._oct_store = dict(prefixes=torch.empty(cap, dtype=kt, device=dev),
                               child_offsets=torch.empty(cap + 1, dtype=i32, device=dev),
                               parents=torch.empty(max(1, cap // 8), dtype=i32, device=dev),
                               level_range=torch.empty(ops.max_level(self.bits) + 2, dtype=i32, device=dev),
                               internal_to_leaf=torch.empty(cap, dtype=i32, device=dev),
                               leaf_to_internal=torch.empty(cap, dtype=i32, device=dev))
        self._centers = torch.empty((cap, 3), dtype=self.fd, device=dev)
        self._sizes = torch.empty((cap, 3), dtype=self.fd, device=dev)
        self._layout = torch.empty(cap + 1, dtype=i32, device=dev)
        self._radii = torch.empty(cap, dtype=torch.float32, device=dev)
        self._flags = torch.empty((self.vranks, cap), dtype=i32, device=dev)

    def step(self, x, y, z, h, box, cold=True, time_stages=False):
        """one pass; x,y,z,h are device tensors of length n (any order).  cold: build the leaf array from the root
        (first sync); otherwise one update step from the previous leaf array (steady-state sync)."""
        n = self.n
        self.events = [] if time_stages else None
        self._mark("start")
        # a1/a2 keys
        ops.compute_sfc_keys(x, y, z, box, self.keys, self.bits, self.kind)
        self._mark("sfc_keys")
        # a4 sort (iota fused)
        order = self.sorter.set_map_from_codes(self.keys)
        self._mark("sort")
        # a5 gather x,y,z,h with one read of the map
        ops.gather_multi(order, [x, y, z, h], self.sorted)
        xs, ys, zs, hs = self.sorted
        self._mark("gather")
        # a6-a9 leaf array
        if cold or self.first_call:
            self.leaves.compute(self.keys, self.bucket)
            self.first_call = False
        else:
            self.leaves.update(self.keys, self.bucket)
        L = self.leaves.num_nodes
        self._mark("leaf_array")
        # a10 linking
        self._alloc_tree_buffers(L)
        I = (L - 1) // 7
        M = L + I
        s = self._oct_store
        v = ops.OctreeView(L, I, M, s["prefixes"], s["child_offsets"], s["parents"], s["level_range"],
                           s["internal_to_leaf"], s["leaf_to_internal"])
        need = call("build_octree_temp_bytes_k%d" % self.bits, L)
        temp = self.ws.get(need)
        call("build_octree_k%d" % self.bits, ops._p(self.leaves.tree_a), L, ops._p(v.prefixes), ops._p(v.child_offsets),
             ops._p(v.parents), ops._p(v.level_range), ops._p(v.internal_to_leaf), ops._p(v.leaf_to_internal),
             ops._p(temp), temp.numel(), ops._stream())
        self.octree = v
        self._mark("link_tree")
        # a11 centres
        ts = "f32" if self.fd == torch.float32 else "f64"
        call("compute_geo_centers_k%d_%s" % (self.bits, ts), self.kind, ops._p(v.prefixes), M, ops._p(self._centers),
             ops._p(self._sizes), C.byref(box), ops._stream())
        self.centers, self.sizes = self._centers, self._sizes
        self._mark("geo_centers")
        # a14 layout: single rank => every leaf is assigned; layout = exclusive scan of counts
        flags0 = self._flags[0]
        ops.compute_node_layout(self.leaves.counts_buf[:L], flags0[:L], 0, L, self._layout[:L + 1], self.ws)
        self.layout = self._layout
        self._mark("layout")
        # a12 radii of all leaves
        ops.halo_radii(hs, self._layout[:L + 1], 2.0, self._radii[:L])
        self.radii = self._radii
        self._mark("halo_radii")
        # a13 halo discovery for `vranks` virtual ranks (equal leaf ranges): every leaf is searched exactly once as a
        # local leaf, which is the total halo-discovery work the ranks of a distributed run do on this tree
        self._flags[:, :L].zero_()
        for r in range(self.vranks):
            first, last = r * L // self.vranks, (r + 1) * L // self.vranks
            call("find_halos_k%d_%s" % (self.bits, ts), self.kind, ops._p(v.prefixes), ops._p(v.child_offsets),
                 ops._p(v.internal_to_leaf), ops._p(self.leaves.tree_a), ops._p(self._radii), C.byref(box), first, last,
                 ops._p(self._flags[r]), ops._stream())
        self.flags = self._flags
        self._mark("find_halos")
        # a15 neighbours
        if self.with_neighbors:
            ops.find_neighbors(xs, ys, zs, hs, 0, n, box, v, self._layout[:L + 1], self._centers, self._sizes, self.ngmax,
                               self.neighbors, self.neighbors_count)
        self._mark("find_neighbors")
        return L

    def stage_ms(self):
        """per-stage device time of the last step(time_stages=True); call after a synchronize"""
        out = {}
        for (_, a), (name, b) in zip(self.events[:-1], self.events[1:]):
            out[name] = a.elapsed_time(b)
        return out
