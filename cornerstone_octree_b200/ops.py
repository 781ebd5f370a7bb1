"""Host-side mirror of the reference's free functions on the hot path, on torch CUDA tensors.

Every function forwards to ONE C-ABI entry point of libcstone_b200.so (include/cstone_b200.h) on torch's current
CUDA stream; torch only provides device memory and streams.  Names follow the reference
(computeSfcKeysGpu -> compute_sfc_keys, sortByKeyGpu -> sort_by_key, computeNodeCountsGpu -> compute_node_counts,
computeNodeOpsGpu -> compute_node_ops, rebalanceTreeGpu -> rebalance_tree, updateOctreeGpu -> update_octree,
buildOctreeGpu -> build_octree, computeGeoCentersGpu -> compute_geo_centers, findHalosGpu -> find_halos,
findNeighbors -> find_neighbors).

Unsigned keys are carried in signed torch tensors of the same width (int32 / int64 bit patterns): torch's
unsigned dtypes lack most operators and the library only sees raw pointers anyway.
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from ._lib import HILBERT, MORTON, BoxF32, BoxF64, call

KEY_TORCH = {32: torch.int32, 64: torch.int64}


def _ks(t):
    return {torch.int32: "k32", torch.int64: "k64"}[t.dtype]


def _ts(t):
    return {torch.float32: "f32", torch.float64: "f64"}[t.dtype]


def key_bits(t):
    return 32 if t.dtype == torch.int32 else 64


def max_level(bits):
    return 10 if bits == 32 else 21


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise ValueError("cornerstone_octree_b200 operates on CUDA tensors only (no CPU fallback)")
        if t is not None and not t.is_contiguous():
            raise ValueError("tensor must be contiguous")


def make_box(limits, boundaries=(0, 0, 0), dtype=torch.float64):
    """POD equivalent of cstone::Box<T> [ref: sfc/box.hpp:111-137]: lengths and inverse lengths are computed once,
    in T, on the host, exactly like the reference constructor does."""
    nd = np.float32 if dtype in (torch.float32, np.float32) else np.float64
    lim = np.asarray(limits, dtype=nd)
    if lim.size == 2:
        lim = np.array([lim[0], lim[1]] * 3, dtype=nd)
    box = BoxF32() if nd == np.float32 else BoxF64()
    for i in range(6):
        box.limits[i] = lim[i]
    for d in range(3):
        length = nd(lim[2 * d + 1] - lim[2 * d])
        box.lengths[d] = length
        box.inv_lengths[d] = nd(nd(1) / length)
        box.boundaries[d] = int(boundaries[d])
    return box


class Workspace:
    """grow-only device scratch buffer handed to the C ABI as `temp` (the library never allocates)"""

    def __init__(self, device="cuda"):
        self.device = device
        self.buf = None

    def get(self, nbytes):
        nbytes = int(nbytes) + 256
        if self.buf is None or self.buf.numel() < nbytes:
            self.buf = torch.empty(int(nbytes * 1.25), dtype=torch.uint8, device=self.device)
        return self.buf


_default_ws = {}


def _ws(device):
    key = str(device)
    if key not in _default_ws:
        _default_ws[key] = Workspace(device)
    return _default_ws[key]


# --------------------------------------------------------------------------------------------- SFC
def compute_sfc_keys(x, y, z, box, keys=None, key_bits_=64, kind=HILBERT):
    """[ref: sfc/sfc_gpu.h:38-39 computeSfcKeysGpu] keys in/out: entries equal to removeKey are preserved"""
    _require_cuda(x, y, z, keys)
    if keys is None:
        keys = torch.zeros(x.numel(), dtype=KEY_TORCH[key_bits_], device=x.device)
    call("compute_sfc_keys_%s_%s" % (_ks(keys), _ts(x)), kind, _p(x), _p(y), _p(z), _p(keys), x.numel(),
         C.byref(box), _stream())
    return keys


def encode_sfc(ix, iy, iz, key_bits_=64, kind=HILBERT):
    keys = torch.empty(ix.numel(), dtype=KEY_TORCH[key_bits_], device=ix.device)
    call("encode_sfc_" + _ks(keys), kind, _p(ix), _p(iy), _p(iz), _p(keys), ix.numel(), _stream())
    return keys


def decode_sfc(keys, kind=HILBERT):
    out = [torch.empty(keys.numel(), dtype=torch.int32, device=keys.device) for _ in range(3)]
    call("decode_sfc_" + _ks(keys), kind, _p(keys), _p(out[0]), _p(out[1]), _p(out[2]), keys.numel(), _stream())
    return out


def minmax3(x, y, z, ws=None):
    """[ref: primitives/primitives_gpu.h:58-62 MinMaxGpu] -> numpy [xmin,xmax,ymin,ymax,zmin,zmax]"""
    _require_cuda(x, y, z)
    ws = ws or _ws(x.device)
    temp = ws.get(call("minmax_temp_bytes"))
    out = np.zeros(6, dtype=np.float32 if x.dtype == torch.float32 else np.float64)
    call("minmax3_" + _ts(x), _p(x), _p(y), _p(z), x.numel(), out.ctypes.data_as(C.c_void_p), _p(temp), _stream())
    return out


# --------------------------------------------------------------------------------------------- sort / gather
class SfcSorter:
    """[ref: primitives/gather.cuh:44-157 GpuSfcSorter] owns the alternate buffers and the scratch of the sort"""

    def __init__(self, device="cuda"):
        self.device = device
        self.key_buf = None
        self.val_buf = None
        self.temp = None
        self.order = None

    def _reserve(self, keys):
        n = keys.numel()
        if self.key_buf is None or self.key_buf.numel() < n or self.key_buf.dtype != keys.dtype:
            self.key_buf = torch.empty(n, dtype=keys.dtype, device=keys.device)
            self.val_buf = torch.empty(n, dtype=torch.int32, device=keys.device)
        need = call("sort_by_key_temp_bytes_" + _ks(keys), n)
        if self.temp is None or self.temp.numel() < need:
            self.temp = torch.empty(int(need), dtype=torch.uint8, device=keys.device)
        return need

    def set_map_from_codes(self, keys):
        """sort `keys` in place; the permutation (initially iota, never materialised) is returned"""
        _require_cuda(keys)
        n = keys.numel()
        need = self._reserve(keys)
        if self.order is None or self.order.numel() != n:
            self.order = torch.empty(n, dtype=torch.int32, device=keys.device)
        call("sort_by_key_" + _ks(keys), _p(keys), _p(self.order), _p(self.key_buf), _p(self.val_buf), n,
             _p(self.temp), need, 1, _stream())
        return self.order

    def update_map(self, keys):
        """re-sort with the existing permutation as payload [ref: primitives/gather.cuh:110-122 updateMap]"""
        n = keys.numel()
        need = self._reserve(keys)
        call("sort_by_key_" + _ks(keys), _p(keys), _p(self.order), _p(self.key_buf), _p(self.val_buf), n,
             _p(self.temp), need, 0, _stream())
        return self.order


def sort_by_key(keys, values=None, sorter=None):
    """[ref: primitives/primitives_gpu.h:104-110 sortByKeyGpu] in place; returns (keys, values)"""
    sorter = sorter or SfcSorter(keys.device)
    if values is None:
        return keys, sorter.set_map_from_codes(keys)
    sorter.order = values
    return keys, sorter.update_map(keys)


def gather(order, src, dst=None):
    """[ref: primitives/primitives_gpu.h:48-49 gatherGpu] dst[i] = src[order[i]]"""
    _require_cuda(order, src, dst)
    n = order.numel()
    if dst is None:
        dst = torch.empty((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    elem = src.element_size() * int(np.prod(src.shape[1:], dtype=np.int64))
    call("gather", _p(order), n, _p(src), _p(dst), elem, _stream())
    return dst


def scatter(order, src, dst):
    """[ref: primitives/primitives_gpu.h:55-56 scatterGpu] dst[order[i]] = src[i]"""
    elem = src.element_size() * int(np.prod(src.shape[1:], dtype=np.int64))
    call("scatter", _p(order), order.numel(), _p(src), _p(dst), elem, _stream())
    return dst


def gather_multi(order, srcs, dsts=None):
    """gather several 4- or 8-byte arrays with one read of the permutation"""
    n = order.numel()
    if dsts is None:
        dsts = [torch.empty(n, dtype=s.dtype, device=s.device) for s in srcs]
    k = len(srcs)
    src_arr = (C.c_void_p * k)(*[s.data_ptr() for s in srcs])
    dst_arr = (C.c_void_p * k)(*[d.data_ptr() for d in dsts])
    eb = (C.c_int * k)(*[s.element_size() for s in srcs])
    call("gather_multi", _p(order), n, k, src_arr, dst_arr, eb, _stream())
    return dsts


def exclusive_scan(values, init=0, out=None, ws=None):
    ws = ws or _ws(values.device)
    out = torch.empty_like(values) if out is None else out
    temp = ws.get(call("scan_temp_bytes", values.numel()))
    call("exclusive_scan_u32", _p(values), _p(out), values.numel(), init, _p(temp), _stream())
    return out


def inclusive_scan(values, out=None, ws=None):
    ws = ws or _ws(values.device)
    out = torch.empty_like(values) if out is None else out
    temp = ws.get(call("scan_temp_bytes", values.numel()))
    call("inclusive_scan_u32", _p(values), _p(out), values.numel(), _p(temp), _stream())
    return out


# --------------------------------------------------------------------------------------------- leaf array
def compute_node_counts(tree, keys, counts=None, max_count=0xFFFFFFFF, use_counts_as_guess=False, num_nodes=None):
    """[ref: tree/csarray_gpu.h:55-62 computeNodeCountsGpu]"""
    _require_cuda(tree, keys, counts)
    n = tree.numel() - 1 if num_nodes is None else num_nodes
    if counts is None:
        counts = torch.empty(n, dtype=torch.int32, device=tree.device)
    first = keys.data_ptr()
    last = first + keys.numel() * keys.element_size()
    call("compute_node_counts_" + _ks(tree), _p(tree), _p(counts), n, C.c_void_p(first), C.c_void_p(last), max_count,
         int(use_counts_as_guess), _stream())
    return counts


def compute_node_ops(tree, counts, bucket_size, node_ops=None, num_nodes=None, ws=None):
    """[ref: tree/csarray_gpu.h:75-77 computeNodeOpsGpu] -> (scanned node_ops, new_num_nodes, converged)"""
    ws = ws or _ws(tree.device)
    n = tree.numel() - 1 if num_nodes is None else num_nodes
    if node_ops is None:
        node_ops = torch.empty(n + 1, dtype=torch.int32, device=tree.device)
    temp = ws.get(call("scan_temp_bytes", n + 1) + 256)
    new_n, conv = C.c_int32(0), C.c_int32(0)
    call("compute_node_ops_" + _ks(tree), _p(tree), n, _p(counts), bucket_size, _p(node_ops), C.byref(new_n),
         C.byref(conv), _p(temp), _stream())
    return node_ops, new_n.value, bool(conv.value)


def rebalance_decision(tree, counts, bucket_size, num_nodes=None):
    n = tree.numel() - 1 if num_nodes is None else num_nodes
    ops = torch.empty(n, dtype=torch.int32, device=tree.device)
    call("rebalance_decision_" + _ks(tree), _p(tree), n, _p(counts), bucket_size, _p(ops), _stream())
    return ops


def rebalance_tree(tree, node_ops, new_num_nodes, new_tree=None, num_nodes=None):
    """[ref: tree/csarray_gpu.h:79-84 rebalanceTreeGpu]"""
    n = tree.numel() - 1 if num_nodes is None else num_nodes
    if new_tree is None:
        new_tree = torch.empty(new_num_nodes + 1, dtype=tree.dtype, device=tree.device)
    call("rebalance_tree_" + _ks(tree), _p(tree), n, new_num_nodes, _p(node_ops), _p(new_tree), _stream())
    return new_tree


class LeafArray:
    """Device-resident cornerstone leaf array with capacity-managed ping-pong buffers; wraps the update / build
    drivers [ref: tree/update_gpu.cuh:58-82 updateOctreeGpu, tree/csarray.hpp:451-465 computeOctree]."""

    def __init__(self, key_bits_=64, capacity=1 << 16, device="cuda"):
        self.kt = KEY_TORCH[key_bits_]
        self.bits = key_bits_
        self.device = device
        self.num_nodes = 0
        self._alloc(capacity)

    def _alloc(self, capacity):
        self.capacity = int(capacity)
        self.tree_a = torch.empty(self.capacity + 1, dtype=self.kt, device=self.device)
        self.tree_b = torch.empty(self.capacity + 1, dtype=self.kt, device=self.device)
        self.counts_buf = torch.empty(self.capacity, dtype=torch.int32, device=self.device)
        self.node_ops = torch.empty(self.capacity + 1, dtype=torch.int32, device=self.device)
        self.temp = torch.empty(int(call("scan_temp_bytes", self.capacity + 1)) + 512, dtype=torch.uint8,
                                device=self.device)

    def _grow(self, needed):
        old = (self.tree_a, self.counts_buf, self.num_nodes)
        self._alloc(int(needed * 1.05) + 64)
        n = old[2]
        if n:
            self.tree_a[:n + 1].copy_(old[0][:n + 1])
            self.counts_buf[:n].copy_(old[1][:n])

    @property
    def tree(self):
        return self.tree_a[:self.num_nodes + 1]

    @property
    def counts(self):
        return self.counts_buf[:self.num_nodes]

    def set(self, tree, counts):
        n = tree.numel() - 1
        if n > self.capacity:
            self._alloc(n)
        self.tree_a[:n + 1].copy_(tree)
        self.counts_buf[:n].copy_(counts)
        self.num_nodes = n

    def _keys_range(self, keys):
        first = keys.data_ptr()
        return C.c_void_p(first), C.c_void_p(first + keys.numel() * keys.element_size())

    def update(self, keys, bucket_size, max_count=0xFFFFFFFF):
        """one rebalance + recount step; returns converged"""
        first, last = self._keys_range(keys)
        while True:
            n, conv = C.c_int32(self.num_nodes), C.c_int32(0)
            st = _lib.lib.__getattr__("cstone_b200_update_octree_" + _ks(self.tree_a))(
                first, last, bucket_size, max_count, _p(self.tree_a), _p(self.tree_b), _p(self.counts_buf),
                _p(self.node_ops), self.capacity, C.byref(n), C.byref(conv), _p(self.temp), _stream())
            if st == -5:  # CSTONE_E_CAPACITY: n holds the required size, inputs untouched
                self._grow(n.value)
                continue
            if st != 0:
                raise _lib.CstoneError(st, "cstone_b200_update_octree")
            self.num_nodes = n.value
            return bool(conv.value)

    def compute(self, keys, bucket_size, max_count=0xFFFFFFFF):
        """build from the root until converged; returns the number of update steps"""
        first, last = self._keys_range(keys)
        while True:
            n, iters = C.c_int32(0), C.c_int32(0)
            st = _lib.lib.__getattr__("cstone_b200_compute_octree_" + _ks(self.tree_a))(
                first, last, bucket_size, max_count, _p(self.tree_a), _p(self.tree_b), _p(self.counts_buf),
                _p(self.node_ops), self.capacity, C.byref(n), C.byref(iters), _p(self.temp), _stream())
            if st == -5:
                self.num_nodes = 0
                self._grow(n.value)
                continue
            if st != 0:
                raise _lib.CstoneError(st, "cstone_b200_compute_octree")
            self.num_nodes = n.value
            return iters.value


# --------------------------------------------------------------------------------------------- internal tree
@dataclass
class OctreeView:
    """[ref: tree/octree.hpp:279-293 OctreeView] device arrays of the fully linked octree"""
    num_leaf_nodes: int
    num_internal_nodes: int
    num_nodes: int
    prefixes: torch.Tensor
    child_offsets: torch.Tensor
    parents: torch.Tensor
    level_range: torch.Tensor
    internal_to_leaf: torch.Tensor
    leaf_to_internal: torch.Tensor


def build_octree(leaves, num_leaf_nodes=None, ws=None):
    """[ref: tree/octree_gpu.h:49-50 buildOctreeGpu; sizes as OctreeData::resize tree/octree.hpp:327-342]"""
    _require_cuda(leaves)
    ws = ws or _ws(leaves.device)
    L = leaves.numel() - 1 if num_leaf_nodes is None else num_leaf_nodes
    I = (L - 1) // 7
    M = L + I
    dev = leaves.device
    i32 = torch.int32
    v = OctreeView(L, I, M, torch.empty(M, dtype=leaves.dtype, device=dev), torch.zeros(M + 1, dtype=i32, device=dev),
                   torch.zeros(max(1, (M - 1) // 8), dtype=i32, device=dev),
                   torch.zeros(max_level(key_bits(leaves)) + 2, dtype=i32, device=dev),
                   torch.empty(M, dtype=i32, device=dev), torch.empty(M, dtype=i32, device=dev))
    need = call("build_octree_temp_bytes_" + _ks(leaves), L)
    temp = ws.get(need)
    call("build_octree_" + _ks(leaves), _p(leaves), L, _p(v.prefixes), _p(v.child_offsets), _p(v.parents),
         _p(v.level_range), _p(v.internal_to_leaf), _p(v.leaf_to_internal), _p(temp), temp.numel(), _stream())
    return v


def upsweep_sum(level_range_host, child_offsets, counts):
    """[ref: tree/octree_gpu.h:52-53 upsweepSumGpu] level_range_host: numpy int32 array (host), counts updated in place"""
    lr = np.ascontiguousarray(level_range_host, dtype=np.int32)
    call("upsweep_sum", lr.size - 1, lr.ctypes.data_as(C.c_void_p), _p(child_offsets), _p(counts), _stream())
    return counts


def compute_geo_centers(prefixes, box, dtype=torch.float64, kind=HILBERT, num_nodes=None):
    """[ref: focus/source_center_gpu.h computeGeoCentersGpu] -> centers[M,3], sizes[M,3]"""
    n = prefixes.numel() if num_nodes is None else num_nodes
    centers = torch.empty((n, 3), dtype=dtype, device=prefixes.device)
    sizes = torch.empty((n, 3), dtype=dtype, device=prefixes.device)
    call("compute_geo_centers_%s_%s" % (_ks(prefixes), _ts(centers)), kind, _p(prefixes), n, _p(centers), _p(sizes),
         C.byref(box), _stream())
    return centers, sizes


# --------------------------------------------------------------------------------------------- halos
def halo_radii(h, segments, scale=2.0, radii=None):
    """[ref: halos/halos.hpp:156-158 segmentMax + scaleGpu] segments: uint32 offsets (int32 tensor), len n+1"""
    n = segments.numel() - 1
    if radii is None:
        radii = torch.empty(n, dtype=torch.float32, device=h.device)
    call("halo_radii_" + _ts(h), _p(h), _p(segments), n, scale, _p(radii), _stream())
    return radii


def find_halos(octree, leaves, radii, box, first_node, last_node, flags=None, kind=HILBERT):
    """[ref: traversal/collisions_gpu.h:58-67 findHalosGpu] flags must be zero-initialised (done here if None)"""
    L = octree.num_leaf_nodes
    if flags is None:
        flags = torch.zeros(L, dtype=torch.int32, device=leaves.device)
    ts = "f32" if isinstance(box, BoxF32) else "f64"
    call("find_halos_%s_%s" % (_ks(leaves), ts), kind, _p(octree.prefixes), _p(octree.child_offsets),
         _p(octree.internal_to_leaf), _p(leaves), _p(radii), C.byref(box), first_node, last_node, _p(flags), _stream())
    return flags


def compute_node_layout(counts, flags, first_assigned, last_assigned, layout=None, ws=None):
    """[ref: domain/layout.hpp:150-164 computeNodeLayout]"""
    ws = ws or _ws(counts.device)
    L = counts.numel()
    if layout is None:
        layout = torch.empty(L + 1, dtype=torch.int32, device=counts.device)
    temp = ws.get(call("scan_temp_bytes", L + 1))
    call("compute_node_layout", _p(counts), _p(flags), first_assigned, last_assigned, L, _p(layout), _p(temp),
         _stream())
    return layout


# --------------------------------------------------------------------------------------------- neighbours
def find_neighbors(x, y, z, h, first_id, last_id, box, octree, layout, centers, sizes, ngmax, neighbors=None,
                   neighbors_count=None, search_ext_factor=1.0, use_workspace=True):
    """[ref: findneighbors.hpp:167-178 findNeighbors] -> neighbors[(last-first), ngmax], neighbors_count"""
    _require_cuda(x, y, z, h, layout, centers, sizes)
    nw = last_id - first_id
    if neighbors is None:
        neighbors = torch.empty((nw, ngmax), dtype=torch.int32, device=x.device)
    if neighbors_count is None:
        neighbors_count = torch.empty(nw, dtype=torch.int32, device=x.device)
    num_leaves = layout.numel() - 1
    temp = _ws(x.device).get(call("find_neighbors_temp_bytes", num_leaves)) if use_workspace else None
    if temp is None:
        call("find_neighbors_" + _ts(x), _p(x), _p(y), _p(z), _p(h), first_id, last_id, C.byref(box),
             _p(octree.child_offsets), _p(octree.internal_to_leaf), _p(layout), _p(centers), _p(sizes),
             search_ext_factor, ngmax, _p(neighbors), _p(neighbors_count), _stream())
    else:
        call("find_neighbors_ws_" + _ts(x), _p(x), _p(y), _p(z), _p(h), first_id, last_id, C.byref(box),
             _p(octree.child_offsets), _p(octree.internal_to_leaf), _p(layout), _p(centers), _p(sizes),
             search_ext_factor, ngmax, _p(neighbors), _p(neighbors_count), num_leaves, _p(temp), _stream())
    return neighbors, neighbors_count


# --------------------------------------------------------------------------------------------- small primitives
# [ref: primitives/primitives_gpu.h:39-125, tree/csarray_gpu.h:86-89, halos/gather_halos_gpu.h:21-27]
_INT_SFX = {torch.int32: "u32", torch.int64: "u64"}


def _host_scalar(ctype, v=0):
    return ctype(v)


def fill(t, value):
    """fillGpu: every element of a 1-, 4- or 8-byte tensor = value"""
    _require_cuda(t)
    v = np.array([value], dtype={torch.uint8: np.uint8, torch.int8: np.int8, torch.int32: np.int32, torch.int64: np.int64,
                                 torch.float32: np.float32, torch.float64: np.float64}[t.dtype])
    call("fill", _p(t), t.numel(), v.ctypes.data_as(C.c_void_p), t.element_size(), _stream())
    return t


def scale(t, value):
    """scaleGpu: t *= value in place"""
    _require_cuda(t)
    call("scale_" + _ts(t), _p(t), t.numel(), value, _stream())
    return t


def increment(src, value, out=None):
    """incrementGpu: out[i] = src[i] + value (wrapping, 32- or 64-bit integers)"""
    _require_cuda(src, out)
    out = torch.empty_like(src) if out is None else out
    bits = 32 if src.dtype == torch.int32 else 64
    call("increment_" + _INT_SFX[src.dtype], _p(src), _p(out), src.numel(), int(value) & ((1 << bits) - 1), _stream())
    return out


def max_norm_square(x, y, z, ws=None):
    _require_cuda(x, y, z)
    ws = ws or _ws(x.device)
    temp = ws.get(call("reduce_temp_bytes"))
    out = np.zeros(1, dtype=np.float32 if x.dtype == torch.float32 else np.float64)
    call("max_norm_square_" + _ts(x), _p(x), _p(y), _p(z), x.numel(), out.ctypes.data_as(C.c_void_p), _p(temp), _stream())
    return out[0]


def reduce_sum(t, init=0, ws=None):
    """reduceGpu<unsigned, size_t>: 64-bit sum of a u32 array (int32 bit patterns)"""
    _require_cuda(t)
    ws = ws or _ws(t.device)
    temp = ws.get(call("reduce_temp_bytes"))
    out = C.c_uint64(0)
    call("reduce_sum_u32", _p(t), t.numel(), int(init), C.byref(out), _p(temp), _stream())
    return int(out.value)


def count(t, value, ws=None):
    """countGpu: number of elements equal to value"""
    _require_cuda(t)
    ws = ws or _ws(t.device)
    temp = ws.get(call("reduce_temp_bytes"))
    out = C.c_uint64(0)
    bits = 32 if t.dtype == torch.int32 else 64
    call("count_" + _INT_SFX[t.dtype], _p(t), t.numel(), int(value) & ((1 << bits) - 1), C.byref(out), _p(temp), _stream())
    return int(out.value)


def lower_bound(sorted_t, value, unsigned=True, ws=None):
    """lowerBoundGpu(first, last, value): index of the first element >= value"""
    _require_cuda(sorted_t)
    ws = ws or _ws(sorted_t.device)
    temp = ws.get(call("reduce_temp_bytes"))
    out = C.c_uint64(0)
    if sorted_t.dtype == torch.float32:
        sfx, v = "f32", float(value)
    else:
        bits = 32 if sorted_t.dtype == torch.int32 else 64
        sfx = ("u" if unsigned else "i") + str(bits)
        v = int(value) & ((1 << bits) - 1) if unsigned else int(value)
    call("lower_bound_" + sfx, _p(sorted_t), sorted_t.numel(), v, C.byref(out), _p(temp), _stream())
    return int(out.value)


def lower_bound_many(sorted_keys, values, index_bits=32):
    """lowerBoundGpu(first, last, valueFirst, valueLast, result) on unsigned keys (signed torch bit patterns)"""
    _require_cuda(sorted_keys, values)
    res = torch.empty(values.numel(), dtype=KEY_TORCH[index_bits], device=values.device)
    call("lower_bound_many_%s_u%d" % (_ks(sorted_keys), index_bits), _p(sorted_keys), sorted_keys.numel(), _p(values),
         values.numel(), _p(res), _stream())
    return res


def segment_max(inp, segments, out_dtype=None):
    """segmentMax<Tin, Tout, IndexType>: segments int32 / int64 (unsigned bit patterns), num_segments+1 entries"""
    _require_cuda(inp, segments)
    out_dtype = out_dtype or inp.dtype
    nseg = segments.numel() - 1
    out = torch.empty(max(nseg, 0), dtype=out_dtype, device=inp.device)
    if nseg > 0:
        call("segment_max_%s_%s" % (_ts(inp), _ts(out)), _p(inp), _p(segments), segments.element_size(), nseg,
             inp.numel(), _p(out), _stream())
    return out


def sort_keys(keys, ws=None):
    """sortGpu(first, last, keyBuf): in-place ascending sort of unsigned keys"""
    _require_cuda(keys)
    ws = ws or _ws(keys.device)
    ks = _ks(keys)
    nbytes = call("sort_keys_temp_bytes_" + ks, keys.numel())
    temp = ws.get(nbytes)
    buf = torch.empty_like(keys)
    call("sort_keys_" + ks, _p(keys), _p(buf), keys.numel(), _p(temp), nbytes, _stream())
    return keys


def count_sfc_gaps(tree):
    """countSfcGapsGpu: node_ops[i] = nodes needed to span [tree[i], tree[i+1]); num_nodes+1 entries, last = 0"""
    _require_cuda(tree)
    n = tree.numel() - 1
    ops_ = torch.zeros(n + 1, dtype=torch.int32, device=tree.device)
    call("count_sfc_gaps_" + _ks(tree), _p(tree), n, _p(ops_), _stream())
    return ops_


def fill_sfc_gaps(tree, scanned_ops, new_num_nodes):
    """fillSfcGapsGpu with the exclusive scan of count_sfc_gaps"""
    _require_cuda(tree, scanned_ops)
    new_tree = torch.empty(new_num_nodes + 1, dtype=tree.dtype, device=tree.device)
    call("fill_sfc_gaps_" + _ks(tree), _p(tree), tree.numel() - 1, _p(scanned_ops), _p(new_tree), _stream())
    return new_tree


def gather_ranges(range_scan, range_offsets, src, buffer_size):
    """gatherRanges<T, IndexType>: pack index ranges of src (rows of 4/8/12/16 bytes) into one send buffer"""
    _require_cuda(range_scan, range_offsets, src)
    row_bytes = src.element_size() * (src.shape[1] if src.dim() == 2 else 1)
    out = torch.empty((buffer_size,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    call("gather_ranges", _p(range_scan), _p(range_offsets), range_scan.element_size(), range_scan.numel(), _p(src),
         _p(out), buffer_size, row_bytes, _stream())
    return out


# --------------------------------------------------------------------------------------------- focus-tree kernels
# [ref: focus/rebalance_gpu.h:39-85, traversal/collisions_gpu.h markMacsGpu, focus/source_center_gpu.h:72-89]
def _key_arg(t, v):
    return int(v) & ((1 << key_bits(t)) - 1)


def rebalance_decision_essential(octree, counts, macs, focus_start, focus_end, bucket_size):
    """rebalanceDecisionEssentialGpu: op per NODE (0 merge / 1 keep / 8 split); macs is an int8 tensor"""
    p = octree.prefixes
    _require_cuda(p, counts, macs)
    n = octree.num_nodes
    node_ops = torch.empty(n, dtype=torch.int32, device=p.device)
    call("rebalance_decision_essential_" + _ks(p), _p(p), _p(octree.child_offsets), _p(octree.parents), _p(counts),
         _p(macs), _key_arg(p, focus_start), _key_arg(p, focus_end), bucket_size, _p(node_ops), n, _stream())
    return node_ops


def mac_refine_decision(octree, macs, focus_first, focus_last):
    p = octree.prefixes
    L = octree.num_leaf_nodes
    node_ops = torch.empty(L, dtype=torch.int32, device=p.device)
    l2i = octree.leaf_to_internal[octree.num_internal_nodes:]  # the OctreeView span [ref: tree/octree.hpp:355]
    call("mac_refine_decision_" + _ks(p), _p(p), _p(macs), _p(l2i), L, focus_first, focus_last,
         _p(node_ops), _stream())
    return node_ops


def protect_ancestors(octree, node_ops, ws=None):
    """protectAncestorsGpu, in place on node_ops -> converged"""
    p = octree.prefixes
    temp = (ws or _ws(p.device)).get(64)
    conv = C.c_int32(0)
    call("protect_ancestors_" + _ks(p), _p(p), _p(octree.parents), _p(node_ops), octree.num_nodes, C.byref(conv),
         _p(temp), _stream())
    return bool(conv.value)


def enforce_keys(forced_keys, octree, node_ops, ws=None):
    """enforceKeysGpu, in place on node_ops -> ResolutionStatus (0 converged, 1 cancelMerge, 2 rebalance, 3 failed)"""
    p = octree.prefixes
    temp = (ws or _ws(p.device)).get(64)
    st = C.c_int32(0)
    call("enforce_keys_" + _ks(p), _p(forced_keys), forced_keys.numel(), _p(p), _p(octree.child_offsets),
         _p(octree.parents), _p(node_ops), C.byref(st), _p(temp), _stream())
    return int(st.value)


def range_count(leaves, counts, leaves_focus, leaves_focus_idx, counts_focus=None):
    """rangeCountGpu"""
    _require_cuda(leaves, counts, leaves_focus, leaves_focus_idx)
    if counts_focus is None:
        counts_focus = torch.zeros(leaves_focus.numel() - 1, dtype=torch.int32, device=leaves.device)
    call("range_count_" + _ks(leaves), _p(leaves), leaves.numel(), _p(counts), _p(leaves_focus), _p(leaves_focus_idx),
         leaves_focus_idx.numel(), _p(counts_focus), _stream())
    return counts_focus


def geo_mac_spheres(prefixes, inv_theta, box, dtype=torch.float64, kind=HILBERT):
    c4 = torch.empty((prefixes.numel(), 4), dtype=dtype, device=prefixes.device)
    call("geo_mac_spheres_%s_%s" % (_ks(prefixes), _ts(c4)), kind, _p(prefixes), prefixes.numel(), _p(c4), inv_theta,
         C.byref(box), _stream())
    return c4


def set_mac(prefixes, mac_spheres, inv_theta, box, kind=HILBERT):
    call("set_mac_%s_%s" % (_ks(prefixes), _ts(mac_spheres)), kind, _p(prefixes), prefixes.numel(), _p(mac_spheres),
         inv_theta, C.byref(box), _stream())
    return mac_spheres


def mark_macs(octree, centers4, box, focus_nodes, limit_source, markings=None, kind=HILBERT):
    """markMacsGpu: focus_nodes = cornerstone keys of the focus leaves (n + 1 entries); markings int8[num_nodes]"""
    p = octree.prefixes
    if markings is None:
        markings = torch.zeros(octree.num_nodes, dtype=torch.int8, device=p.device)
    call("mark_macs_%s_%s" % (_ks(p), _ts(centers4)), kind, _p(p), _p(octree.child_offsets), _p(centers4), C.byref(box),
         _p(focus_nodes), focus_nodes.numel() - 1, int(limit_source), _p(markings), _stream())
    return markings


def move_centers(src3):
    dst = torch.empty((src3.shape[0], 4), dtype=src3.dtype, device=src3.device)
    call("move_centers_" + _ts(src3), _p(src3), src3.shape[0], _p(dst), _stream())
    return dst


# --------------------------------------------------------------------------------------------- expansion centres
def compute_leaf_source_center(x, y, z, m, octree, layout, centers4=None):
    """computeLeafSourceCenterGpu [ref: focus/source_center_gpu.h:51-66]: centers4[leaf node] = (mass centre, mass)"""
    _require_cuda(x, y, z, m, layout)
    tf = x.dtype
    if centers4 is None:
        centers4 = torch.zeros((octree.num_nodes, 4), dtype=tf, device=x.device)
    l2i = octree.leaf_to_internal[octree.num_internal_nodes:]
    call("compute_leaf_source_center_%s_%s_%s" % (_ts(x), _ts(m), _ts(centers4)), _p(x), _p(y), _p(z), _p(m), _p(l2i),
         octree.num_leaf_nodes, _p(layout), _p(centers4), _stream())
    return centers4


def upsweep_centers(level_range_host, child_offsets, centers4):
    """upsweepCentersGpu [ref: focus/source_center_gpu.h:68-74], in place; level_range_host = numpy int32 (host)"""
    lr = np.ascontiguousarray(level_range_host, dtype=np.int32)
    call("upsweep_centers_" + _ts(centers4), lr.size - 1, lr.ctypes.data_as(C.c_void_p), _p(child_offsets),
         _p(centers4), _stream())
    return centers4


# --------------------------------------------------------------------------------------------- target groups
def compute_fixed_groups(first, last, group_size, device="cuda"):
    """computeFixedGroups [ref: traversal/groups_gpu.h:60]: group starts, last entry = last"""
    n = (last - first + group_size - 1) // group_size
    groups = torch.empty(n + 1, dtype=torch.int32, device=device)
    call("compute_fixed_groups", first, last, group_size, _p(groups), _stream())
    return groups


def compute_group_splits(first, last, x, y, z, h, leaves, layout, box, group_size, tol_factor, ws=None):
    """computeGroupSplits [ref: traversal/groups_gpu.h:75-88] -> int32 tensor of num_groups + 1 group starts"""
    _require_cuda(x, y, z, leaves, layout)
    n = max(last - first, 0)
    tb = int(call("group_splits_temp_bytes", n, group_size))
    temp = (ws or _ws(x.device)).get(max(tb, 16))
    hs = _ts(h) if h is not None else _ts(x)
    groups = torch.empty(n + 1, dtype=torch.int32, device=x.device)
    ng = C.c_uint32(0)
    call("compute_group_splits_%s_%s_%s" % (_ks(leaves), _ts(x), hs), first, last, _p(x), _p(y), _p(z),
         _p(h) if h is not None else None, _p(leaves), leaves.numel() - 1, _p(layout), C.byref(box), group_size,
         tol_factor, _p(groups), groups.numel(), C.byref(ng), _p(temp), temp.numel(), _stream())
    return groups[:ng.value + 1]
