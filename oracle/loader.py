"""TEST INFRASTRUCTURE ONLY — ctypes/numpy access to the CPU checkers (oracle_api.h).

`load("port")`  -> oracle/liboracle_port.so        (our CPU restatement, prefix cstone_oracle_)
`load("ref")`   -> oracle/_ref/liboracle_ref.so    (reference headers compiled here, prefix cstone_ref_)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
module.  The product package (cornerstone_octree_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATHS = {"port": os.path.join(_HERE, "liboracle_port.so"), "ref": os.path.join(_HERE, "_ref", "liboracle_ref.so")}
_PREFIX = {"port": "cstone_oracle_", "ref": "cstone_ref_"}

MORTON, HILBERT = 0, 1
OPEN, PERIODIC, FIXED = 0, 1, 2


class BoxF32(C.Structure):
    _fields_ = [("limits", C.c_float * 6), ("lengths", C.c_float * 3), ("inv_lengths", C.c_float * 3),
                ("boundaries", C.c_ubyte * 3)]


class BoxF64(C.Structure):
    _fields_ = [("limits", C.c_double * 6), ("lengths", C.c_double * 3), ("inv_lengths", C.c_double * 3),
                ("boundaries", C.c_ubyte * 3)]


def make_box(limits, boundaries=(OPEN, OPEN, OPEN), dtype=np.float64):
    """Same arithmetic as cstone::Box<T> ctor [ref: sfc/box.hpp:128-137]: lengths = max-min and
    inverse lengths = T(1)/(max-min), both evaluated in T."""
    dtype = np.dtype(dtype)
    lim = np.asarray(limits, dtype=dtype)
    if lim.size == 2:
        lim = np.array([lim[0], lim[1]] * 3, dtype=dtype)
    box = BoxF32() if dtype == np.float32 else BoxF64()
    one = dtype.type(1)
    for i in range(6):
        box.limits[i] = lim[i]
    for d in range(3):
        length = dtype.type(lim[2 * d + 1] - lim[2 * d])
        box.lengths[d] = length
        box.inv_lengths[d] = dtype.type(one / length)
    for d in range(3):
        box.boundaries[d] = int(boundaries[d])
    return box


def available(which):
    return os.path.exists(_PATHS[which])


def build(which="port"):
    subprocess.check_call(["make", "-C", _HERE, which], stdout=subprocess.DEVNULL)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _ks(dtype):
    return {np.dtype(np.uint32): "k32", np.dtype(np.uint64): "k64"}[np.dtype(dtype)]


def _ts(dtype):
    return {np.dtype(np.float32): "f32", np.dtype(np.float64): "f64"}[np.dtype(dtype)]


def max_level(dtype):
    return 10 if np.dtype(dtype) == np.uint32 else 21


class Oracle:
    """numpy front-end; every method allocates and returns its outputs."""

    def __init__(self, which):
        if not available(which):
            raise FileNotFoundError(_PATHS[which])
        self.which = which
        self.lib = C.CDLL(_PATHS[which])
        self.p = _PREFIX[which]

    def _f(self, name, restype=None):
        f = getattr(self.lib, self.p + name)
        f.restype = restype
        return f

    # ---- misc
    def num_threads(self):
        return self._f("num_threads", C.c_int)()

    def set_num_threads(self, n):
        self._f("set_num_threads")(C.c_int(n))

    # ---- SFC
    def encode(self, kind, ix, iy, iz, kdtype=np.uint64):
        f = self._f("encode_" + _ks(kdtype), C.c_uint32 if np.dtype(kdtype) == np.uint32 else C.c_uint64)
        return f(C.c_int(kind), C.c_uint(ix), C.c_uint(iy), C.c_uint(iz))

    def decode(self, kind, key, kdtype=np.uint64):
        xyz = (C.c_uint * 3)()
        kt = C.c_uint32 if np.dtype(kdtype) == np.uint32 else C.c_uint64
        self._f("decode_" + _ks(kdtype))(C.c_int(kind), kt(int(key)), xyz)
        return tuple(xyz)

    def sfc_ibox(self, kind, key, level, kdtype=np.uint64):
        out = (C.c_int * 6)()
        kt = C.c_uint32 if np.dtype(kdtype) == np.uint32 else C.c_uint64
        self._f("sfc_ibox_" + _ks(kdtype))(C.c_int(kind), kt(int(key)), C.c_uint(level), out)
        return tuple(out)

    def compute_sfc_keys(self, kind, x, y, z, box, kdtype=np.uint64, keys=None):
        n = x.size
        if keys is None:
            keys = np.zeros(n, dtype=kdtype)
        self._f("compute_sfc_keys_%s_%s" % (_ks(keys.dtype), _ts(x.dtype)))(
            C.c_int(kind), _ptr(x), _ptr(y), _ptr(z), _ptr(keys), C.c_size_t(n), C.byref(box))
        return keys

    # ---- sort / gather
    def sort_by_key(self, keys, values=None):
        keys = keys.copy()
        values = np.arange(keys.size, dtype=np.uint32) if values is None else values.copy()
        self._f("sort_by_key_" + _ks(keys.dtype))(_ptr(keys), _ptr(values), C.c_size_t(keys.size))
        return keys, values

    def gather(self, order, src):
        src = np.ascontiguousarray(src)
        dst = np.empty((order.size,) + src.shape[1:], dtype=src.dtype)
        elem = src.dtype.itemsize * int(np.prod(src.shape[1:], dtype=np.int64))
        self._f("gather")(_ptr(order), C.c_size_t(order.size), _ptr(src), _ptr(dst), C.c_int(elem))
        return dst

    # ---- leaf array
    def compute_node_counts(self, tree, keys, max_count=0xFFFFFFFF, guess=False, counts=None):
        n = tree.size - 1
        if counts is None:
            counts = np.zeros(n, dtype=np.uint32)
        first = keys.ctypes.data
        last = first + keys.size * keys.itemsize
        self._f("compute_node_counts_" + _ks(tree.dtype))(
            _ptr(tree), _ptr(counts), C.c_int32(n), C.c_void_p(first), C.c_void_p(last), C.c_uint32(max_count),
            C.c_int(int(guess)))
        return counts

    def rebalance_decision(self, tree, counts, bucket):
        n = tree.size - 1
        ops = np.zeros(n + 1, dtype=np.int32)
        conv = self._f("rebalance_decision_" + _ks(tree.dtype), C.c_int)(
            _ptr(tree), _ptr(counts), C.c_int32(n), C.c_uint32(bucket), _ptr(ops))
        return ops, bool(conv)

    def rebalance_tree(self, tree, ops):
        """ops: n+1 un-scanned entries. returns (new_tree, scanned ops)."""
        n = tree.size - 1
        ops = ops.astype(np.int32).copy()
        cap = int(ops[:n].sum())
        new_tree = np.zeros(cap + 1, dtype=tree.dtype)
        m = self._f("rebalance_tree_" + _ks(tree.dtype), C.c_int32)(
            _ptr(tree), C.c_int32(n), _ptr(ops), _ptr(new_tree), C.c_int32(cap))
        assert m == cap
        return new_tree, ops

    def update_octree(self, keys, bucket, tree, counts, max_count=0xFFFFFFFF):
        n = tree.size - 1
        cap = max(8 * n + 64, 4096 * 8) if n < 64 else 8 * n + 64
        while True:
            t = np.zeros(cap + 1, dtype=tree.dtype)
            c = np.zeros(cap, dtype=np.uint32)
            t[:n + 1] = tree
            c[:n] = counts
            nn = C.c_int32(n)
            first = keys.ctypes.data
            last = first + keys.size * keys.itemsize
            conv = self._f("update_octree_" + _ks(tree.dtype), C.c_int)(
                C.c_void_p(first), C.c_void_p(last), C.c_uint32(bucket), C.c_uint32(max_count), _ptr(t), _ptr(c),
                C.c_int32(cap), C.byref(nn))
            if conv >= 0:
                return t[:nn.value + 1].copy(), c[:nn.value].copy(), bool(conv)
            cap = nn.value

    def compute_octree(self, keys, bucket, max_count=0xFFFFFFFF):
        cap = max(1024, 16 * (keys.size // max(bucket, 1) + 1))
        first = keys.ctypes.data
        last = first + keys.size * keys.itemsize
        while True:
            t = np.zeros(cap + 1, dtype=keys.dtype)
            c = np.zeros(cap, dtype=np.uint32)
            n = self._f("compute_octree_" + _ks(keys.dtype), C.c_int32)(
                C.c_void_p(first), C.c_void_p(last), C.c_uint32(bucket), C.c_uint32(max_count), _ptr(t), _ptr(c),
                C.c_int32(cap))
            if n >= 0:
                return t[:n + 1].copy(), c[:n].copy()
            cap = -n

    # ---- internal tree
    def build_octree(self, leaves):
        L = leaves.size - 1
        I = (L - 1) // 7
        M = L + I
        out = dict(
            prefixes=np.zeros(M, dtype=leaves.dtype), child_offsets=np.zeros(M + 1, dtype=np.int32),
            parents=np.zeros(max(1, (M - 1) // 8), dtype=np.int32),
            level_range=np.zeros(max_level(leaves.dtype) + 2, dtype=np.int32),
            internal_to_leaf=np.zeros(M, dtype=np.int32), leaf_to_internal=np.zeros(M, dtype=np.int32))
        self._f("build_octree_" + _ks(leaves.dtype))(
            _ptr(leaves), C.c_int32(L), _ptr(out["prefixes"]), _ptr(out["child_offsets"]), _ptr(out["parents"]),
            _ptr(out["level_range"]), _ptr(out["internal_to_leaf"]), _ptr(out["leaf_to_internal"]))
        out["num_leaf_nodes"], out["num_internal_nodes"], out["num_nodes"] = L, I, M
        return out

    def span_sfc_range(self, a, b, kdtype=np.uint64):
        """keys of the fewest cornerstone nodes spanning [a, b) [ref: sfc/common.hpp:392-438]"""
        kt = C.c_uint32 if np.dtype(kdtype) == np.uint32 else C.c_uint64
        f = self._f("span_sfc_range_" + _ks(kdtype), C.c_int)
        n = f(kt(int(a)), kt(int(b)), None)
        out = np.zeros(n, dtype=kdtype)
        if n:
            f(kt(int(a)), kt(int(b)), _ptr(out))
        return out

    # ---- focus-tree update functions (reference build only: oracle/ref_wrap.cpp, prefix cstone_ref_)
    def rebalance_decision_essential(self, oc, counts, macs, focus_start, focus_end, bucket):
        """[ref: focus/rebalance.hpp:140-166]"""
        p = oc["prefixes"]
        kt = C.c_uint32 if p.dtype == np.uint32 else C.c_uint64
        ops = np.zeros(p.size, dtype=np.int32)
        self._f("rebalance_decision_essential_" + _ks(p.dtype))(
            _ptr(p), _ptr(oc["child_offsets"]), _ptr(oc["parents"]), _ptr(counts), _ptr(macs), kt(int(focus_start)),
            kt(int(focus_end)), C.c_uint32(bucket), _ptr(ops), C.c_int32(p.size))
        return ops

    def mac_refine_decision(self, oc, macs, focus_first, focus_last):
        p = oc["prefixes"]
        L = oc["num_leaf_nodes"]
        ops = np.zeros(L, dtype=np.int32)
        self._f("mac_refine_decision_" + _ks(p.dtype))(_ptr(p), _ptr(macs), _ptr(np.ascontiguousarray(oc["leaf_to_internal"][oc["num_internal_nodes"]:])), C.c_int32(L),
                                                       C.c_int32(focus_first), C.c_int32(focus_last), _ptr(ops))
        return ops

    def protect_ancestors(self, oc, ops):
        """[ref: focus/rebalance.hpp:168-181] -> (nodeOps, converged)"""
        p = oc["prefixes"]
        ops = ops.copy()
        conv = self._f("protect_ancestors_" + _ks(p.dtype), C.c_int)(_ptr(p), _ptr(oc["parents"]), _ptr(ops),
                                                                     C.c_int32(p.size))
        return ops, bool(conv)

    def enforce_keys(self, keys, oc, ops):
        """[ref: focus/rebalance.hpp:252-266] -> (nodeOps, ResolutionStatus as int)"""
        p = oc["prefixes"]
        ops = ops.copy()
        st = self._f("enforce_keys_" + _ks(p.dtype), C.c_int)(_ptr(keys), C.c_int32(keys.size), _ptr(p),
                                                              _ptr(oc["child_offsets"]), _ptr(oc["parents"]), _ptr(ops))
        return ops, int(st)

    def range_count(self, leaves, counts, leaves_focus, idx):
        """[ref: focus/rebalance.hpp:277-299]"""
        out = np.zeros(leaves_focus.size - 1, dtype=np.uint32)
        self._f("range_count_" + _ks(leaves.dtype))(_ptr(leaves), C.c_int32(leaves.size), _ptr(counts), _ptr(leaves_focus),
                                                    C.c_int32(leaves_focus.size), _ptr(idx), C.c_int32(idx.size), _ptr(out))
        return out

    def geo_mac_spheres(self, prefixes, inv_theta, box, fdtype=np.float64):
        """[ref: traversal/macs.hpp:58-71 computeMinMacR2 per node]"""
        c4 = np.zeros((prefixes.size, 4), dtype=fdtype)
        self._f("geo_mac_spheres_%s_%s" % (_ks(prefixes.dtype), _ts(fdtype)))(
            _ptr(prefixes), C.c_int32(prefixes.size), _ptr(c4), C.c_float(inv_theta), C.byref(box))
        return c4

    def set_mac(self, prefixes, spheres, inv_theta, box):
        """[ref: focus/source_center_gpu.cu setMacKernel over traversal/macs.hpp:84-100 computeVecMacR2]"""
        c4 = spheres.copy()
        self._f("set_mac_%s_%s" % (_ks(prefixes.dtype), _ts(c4.dtype)))(
            _ptr(prefixes), C.c_int32(prefixes.size), _ptr(c4), C.c_float(inv_theta), C.byref(box))
        return c4

    def mark_macs(self, oc, centers4, box, focus_nodes, limit_source):
        """[ref: traversal/macs.hpp:236-269 markMacs]"""
        p = oc["prefixes"]
        marks = np.zeros(p.size, dtype=np.int8)
        self._f("mark_macs_%s_%s" % (_ks(p.dtype), _ts(centers4.dtype)))(
            _ptr(p), _ptr(oc["child_offsets"]), _ptr(centers4), C.byref(box), _ptr(focus_nodes),
            C.c_int32(focus_nodes.size - 1), C.c_int(int(limit_source)), _ptr(marks))
        return marks

    def compute_leaf_source_center(self, x, y, z, m, oc, layout, fdtype=None):
        """[ref: focus/source_center.hpp:110-126 computeLeafMassCenter] (reference build only)"""
        fdtype = np.dtype(fdtype or x.dtype)
        c4 = np.zeros((oc["num_nodes"], 4), dtype=fdtype)
        l2i = np.ascontiguousarray(oc["leaf_to_internal"][oc["num_internal_nodes"]:])
        self._f("compute_leaf_source_center_%s_%s_%s" % (_ts(x.dtype), _ts(m.dtype), _ts(fdtype)))(
            _ptr(x), _ptr(y), _ptr(z), _ptr(m), C.c_size_t(x.size), _ptr(l2i), C.c_int32(l2i.size), _ptr(layout),
            _ptr(c4))
        return c4

    def upsweep_centers(self, level_range, child_offsets, centers4):
        """[ref: tree/octree.hpp:583-601 upsweep with focus/source_center.hpp:81-96 CombineSourceCenter]"""
        c4 = centers4.copy()
        lr = np.ascontiguousarray(level_range, dtype=np.int32)
        self._f("upsweep_centers_" + _ts(c4.dtype))(C.c_int32(lr.size - 1), _ptr(lr), _ptr(child_offsets), _ptr(c4))
        return c4

    def mac_refine_step(self, leaves, old_fs, old_fe, fs, fe, inv_theta, box):
        """one step of macRefine [ref: focus/octree_focus.hpp:260-295] -> (new leaves, converged)"""
        kt = C.c_uint32 if leaves.dtype == np.uint32 else C.c_uint64
        cap = 8 * (leaves.size - 1) + 8
        buf = np.zeros(cap + 1, dtype=leaves.dtype)
        buf[:leaves.size] = leaves
        n = C.c_int32(leaves.size - 1)
        ts = "f32" if isinstance(box, BoxF32) else "f64"
        r = self._f("mac_refine_step_%s_%s" % (_ks(leaves.dtype), ts), C.c_int)(
            _ptr(buf), C.byref(n), C.c_int32(cap), kt(int(old_fs)), kt(int(old_fe)), kt(int(fs)), kt(int(fe)),
            C.c_float(inv_theta), C.byref(box))
        assert r >= 0
        return buf[:n.value + 1].copy(), bool(r)

    def upsweep_sum(self, level_range, child_offsets, counts):
        counts = counts.copy()
        self._f("upsweep_sum")(C.c_int32(level_range.size - 1), _ptr(level_range), _ptr(child_offsets), _ptr(counts))
        return counts

    def geo_centers(self, kind, prefixes, box, fdtype=np.float64):
        n = prefixes.size
        centers = np.zeros((n, 3), dtype=fdtype)
        sizes = np.zeros((n, 3), dtype=fdtype)
        self._f("geo_centers_%s_%s" % (_ks(prefixes.dtype), _ts(fdtype)))(
            C.c_int(kind), _ptr(prefixes), C.c_int32(n), _ptr(centers), _ptr(sizes), C.byref(box))
        return centers, sizes

    # ---- halos
    def halo_radii(self, h, segments, scale=2.0):
        n = segments.size - 1
        radii = np.zeros(n, dtype=np.float32)
        self._f("halo_radii_" + _ts(h.dtype))(_ptr(h), _ptr(segments), C.c_size_t(n), C.c_float(scale), _ptr(radii))
        return radii

    def find_halos(self, kind, octree, leaves, radii, box, first, last, fdtype=np.float64):
        flags = np.zeros(leaves.size - 1, dtype=np.int32)
        self._f("find_halos_%s_%s" % (_ks(leaves.dtype), _ts(fdtype)))(
            C.c_int(kind), _ptr(octree["prefixes"]), _ptr(octree["child_offsets"]), _ptr(octree["internal_to_leaf"]),
            _ptr(leaves), _ptr(radii), C.byref(box), C.c_int32(first), C.c_int32(last), _ptr(flags))
        return flags

    def compute_node_layout(self, counts, flags, first, last):
        L = counts.size
        layout = np.zeros(L + 1, dtype=np.uint32)
        self._f("compute_node_layout")(_ptr(counts), _ptr(flags), C.c_int32(first), C.c_int32(last), C.c_int32(L),
                                       _ptr(layout))
        return layout

    # ---- neighbours
    def find_neighbors(self, x, y, z, h, first, last, box, octree, layout, centers, sizes, ngmax, ext=1.0):
        nw = last - first
        nb = np.zeros(nw * ngmax, dtype=np.uint32)
        nc = np.zeros(nw, dtype=np.uint32)
        self._f("find_neighbors_" + _ts(x.dtype))(
            _ptr(x), _ptr(y), _ptr(z), _ptr(h), C.c_uint32(first), C.c_uint32(last), C.byref(box),
            _ptr(octree["child_offsets"]), _ptr(octree["internal_to_leaf"]), _ptr(layout), _ptr(centers), _ptr(sizes),
            C.c_float(ext), C.c_uint32(ngmax), _ptr(nb), _ptr(nc))
        return nb.reshape(nw, ngmax), nc


_cache = {}


def load(which="port"):
    if which not in _cache:
        _cache[which] = Oracle(which)
    return _cache[which]


class RefFocusTree:
    """the reference's FocusedOctreeSingleNode<KeyType> [ref: focus/octree_focus.hpp:349-427], compiled from its own
    headers into oracle/_ref (reference build only)"""

    def __init__(self, oracle, bucket_size, theta, kdtype=np.uint64):
        assert oracle.which == "ref"
        self.o = oracle
        self.kd = np.dtype(kdtype)
        self.ks = _ks(kdtype)
        self.h = C.c_void_p(oracle._f("focus_tree_create_" + self.ks, C.c_void_p)(C.c_uint32(bucket_size),
                                                                                  C.c_float(theta)))

    def close(self):
        if self.h:
            self.o._f("focus_tree_destroy_" + self.ks)(self.h)
            self.h = None

    def update(self, box, keys, focus_start, focus_end, mandatory=()):
        kt = C.c_uint32 if self.kd == np.uint32 else C.c_uint64
        mand = np.ascontiguousarray(np.asarray(list(mandatory), dtype=self.kd))
        ts = "f32" if isinstance(box, BoxF32) else "f64"
        r = self.o._f("focus_tree_update_%s_%s" % (self.ks, ts), C.c_int)(
            self.h, C.byref(box), _ptr(keys), C.c_size_t(keys.size), kt(int(focus_start)), kt(int(focus_end)),
            _ptr(mand), C.c_int32(mand.size))
        return bool(r)

    def leaves_and_counts(self):
        n = self.o._f("focus_tree_num_leaves_" + self.ks, C.c_int32)(self.h)
        leaves = np.zeros(n + 1, dtype=self.kd)
        counts = np.zeros(n, dtype=np.uint32)
        self.o._f("focus_tree_get_" + self.ks)(self.h, _ptr(leaves), _ptr(counts))
        return leaves, counts


# ------------------------------------------------------------------------------------------------
# multi-rank oracle: the reference's Domain<KeyType,T,CpuTag> on P thread-ranks (oracle/ref_domain.cpp)
# ------------------------------------------------------------------------------------------------
_REFDOM_PATH = os.path.join(_HERE, "_ref", "liboracle_refdom.so")


def refdom_available():
    return os.path.exists(_REFDOM_PATH)


class RefDomain:
    """P ranks of the UNMODIFIED reference Domain (domain/domain.hpp:66-699) in this process, one thread per rank over
    the mpi.h shim.  Drives them as test/integration_mpi/domain_nranks.cpp:103-149 does."""

    def __init__(self, nranks, bucket, bucket_focus, theta, limits, pbc=(0, 0, 0), kdtype=np.uint64, fdtype=np.float64,
                 nprops=0, omp_per_rank=None):
        self.lib = C.CDLL(_REFDOM_PATH)
        self.P, self.kd, self.fd, self.nprops = nranks, np.dtype(kdtype), np.dtype(fdtype), nprops
        self.sfx = "_%s_%s" % (_ks(kdtype), _ts(fdtype))
        lim = np.asarray(limits, dtype=self.fd)
        if lim.size == 2:
            lim = np.array([lim[0], lim[1]] * 3, dtype=self.fd)
        p = np.asarray(pbc, dtype=np.int32)
        if omp_per_rank is None:
            omp_per_rank = max(1, (os.cpu_count() or 1) // nranks)
        f = self._f("create", C.c_void_p)
        self.h = C.c_void_p(f(nranks, C.c_uint(bucket), C.c_uint(bucket_focus), C.c_float(theta), _ptr(lim), _ptr(p),
                              nprops, omp_per_rank))

    def _f(self, name, restype=C.c_int):
        fn = getattr(self.lib, "cstone_refdom_" + name + self.sfx)
        fn.restype = restype
        return fn

    def close(self):
        if self.h:
            self._f("destroy", None)(self.h)
            self.h = None

    def set(self, rank, x, y, z, h, props=()):
        a = [np.ascontiguousarray(v, dtype=self.fd) for v in (x, y, z, h)]
        pr = [np.ascontiguousarray(v, dtype=self.fd) for v in props] + [None, None]
        self._f("set")(self.h, rank, C.c_uint64(a[0].size), *[_ptr(v) for v in a], _ptr(pr[0]), _ptr(pr[1]))

    def sync(self):
        st = self._f("sync")(self.h)
        assert st == 0, "reference Domain::sync failed"

    def set_mass(self, rank, m):
        m = np.ascontiguousarray(m, dtype=self.fd)
        self._f("set_mass")(self.h, rank, C.c_uint64(m.size), _ptr(m))

    def sync_grav(self):
        """Domain::syncGrav on every rank (masses from set_mass)"""
        st = self._f("sync_grav")(self.h)
        assert st == 0, "reference Domain::syncGrav failed"

    def masses(self, rank):
        out = np.zeros(self.info(rank)["n"], dtype=self.fd)
        self._f("get_array")(self.h, rank, 7, _ptr(out))
        return out

    def centers(self, rank):
        """expansion centres of the focus tree after sync_grav: [numNodes, 4] = x, y, z, MAC radius squared"""
        n = np.zeros(1, dtype=np.int64)
        self._f("get_tree")(self.h, rank, 6, _ptr(n))
        out = np.zeros((int(n[0]), 4), dtype=self.fd)
        self._f("get_tree")(self.h, rank, 5, _ptr(out))
        return out

    def info(self, rank):
        out = np.zeros(8, dtype=np.int64)
        box = np.zeros(6, dtype=self.fd)
        self._f("info", None)(self.h, rank, _ptr(out), _ptr(box))
        return dict(start=int(out[0]), end=int(out[1]), size=int(out[2]), focus_leaves=int(out[3]),
                    global_leaves=int(out[4]), start_cell=int(out[5]), end_cell=int(out[6]), n=int(out[7]), box=box)

    def arrays(self, rank):
        """keys, x, y, z, h (+ properties) of one rank after sync, halos included"""
        n = self.info(rank)["n"]
        out = [np.zeros(n, dtype=self.kd)] + [np.zeros(n, dtype=self.fd) for _ in range(4 + self.nprops)]
        for which, a in enumerate(out[:5]):
            self._f("get_array")(self.h, rank, which, _ptr(a))
        for k in range(self.nprops):
            self._f("get_array")(self.h, rank, 5 + k, _ptr(out[5 + k]))
        return out

    def trees(self, rank):
        inf = self.info(rank)
        gl = np.zeros(inf["global_leaves"] + 1, dtype=self.kd)
        fl = np.zeros(inf["focus_leaves"] + 1, dtype=self.kd)
        fc = np.zeros(inf["focus_leaves"], dtype=np.uint32)
        lay = np.zeros(inf["focus_leaves"] + 1, dtype=np.uint32)
        asg = np.zeros(2 * self.P, dtype=np.int32)
        for which, a in enumerate((gl, fl, fc, lay, asg)):
            self._f("get_tree")(self.h, rank, which, _ptr(a))
        return dict(global_leaves=gl, focus_leaves=fl, focus_counts=fc, layout=lay, focus_assignment=asg.reshape(-1, 2))

    def find_neighbors(self, ngmax):
        """the reference's findNeighbors over every rank's assigned range: list of (neighbors[n, ngmax], counts[n])"""
        res, nbp, ncp = [], (C.c_void_p * self.P)(), (C.c_void_p * self.P)()
        for r in range(self.P):
            inf = self.info(r)
            n = inf["end"] - inf["start"]
            nb = np.zeros((n, ngmax), dtype=np.uint32)
            nc = np.zeros(n, dtype=np.uint32)
            res.append((nb, nc))
            nbp[r], ncp[r] = nb.ctypes.data, nc.ctypes.data
        st = self._f("find_neighbors")(self.h, C.c_uint(ngmax), nbp, ncp)
        assert st == 0
        return res

    def exchange_halos(self, arrays):
        """arrays[rank]: fdtype[nParticlesWithHalos(rank)] with the assigned range filled; halos are filled in place"""
        ptrs = (C.c_void_p * self.P)()
        for r in range(self.P):
            assert arrays[r].dtype == self.fd and arrays[r].size == self.info(r)["size"]
            ptrs[r] = arrays[r].ctypes.data
        st = self._f("exchange_halos")(self.h, ptrs)
        assert st == 0

    def bench_step(self, amp, ngmax):
        """bench.py's reference arm: drift (signed amplitude) + Domain::sync + findNeighbors on every rank;
        returns (assigned particles, sum of neighbour counts, halos), each summed over the ranks"""
        out = np.zeros(3, dtype=np.float64)
        st = self._f("bench_step")(self.h, C.c_double(amp), C.c_uint(ngmax), _ptr(out))
        assert st == 0, "reference bench step failed"
        return int(out[0]), float(out[1]), int(out[2])

    def bench_neighbors(self, rank, ngmax):
        """neighbour lists and counts of the last bench_step of one rank (copies)"""
        inf = self.info(rank)
        n = inf["end"] - inf["start"]
        nb, nc = C.POINTER(C.c_uint)(), C.POINTER(C.c_uint)()
        self._f("bench_neighbors", None)(self.h, rank, C.byref(nb), C.byref(nc))
        return (np.ctypeslib.as_array(nb, (n, ngmax)).copy(), np.ctypeslib.as_array(nc, (n,)).copy())

    def seconds(self):
        a, b = C.c_double(0), C.c_double(0)
        self._f("seconds", None)(self.h, C.byref(a), C.byref(b))
        return a.value, b.value
