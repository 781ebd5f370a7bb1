"""numpy-in / numpy-out adapter around cornerstone_octree_b200.ops with the method set of oracle.loader.Oracle,
so that the known-answer tests (kat.py) and the stage-by-stage pipeline comparison (pipeline.py) run unchanged
against the CUDA library.  Every method goes through the C ABI of libcstone_b200.so."""
import numpy as np
import torch

from cornerstone_octree_b200 import ops
from cornerstone_octree_b200._lib import BoxF32, BoxF64

DEV = "cuda"


def _signed(a):
    return a.view({4: np.int32, 8: np.int64}[a.dtype.itemsize]) if a.dtype.kind == "u" else a


def to_dev(a):
    return torch.from_numpy(np.ascontiguousarray(_signed(np.asarray(a)))).to(DEV)


def keys_np(t, kd):
    return t.cpu().numpy().view(np.dtype(kd))


def conv_box(box):
    """oracle.loader box -> product box struct (identical layout)"""
    out = BoxF32() if box.__class__.__name__ == "BoxF32" else BoxF64()
    for i in range(6):
        out.limits[i] = box.limits[i]
    for d in range(3):
        out.lengths[d] = box.lengths[d]
        out.inv_lengths[d] = box.inv_lengths[d]
        out.boundaries[d] = box.boundaries[d]
    return out


class View:
    pass


def oct_to_dev(o):
    v = ops.OctreeView(o["num_leaf_nodes"], o["num_internal_nodes"], o["num_nodes"], to_dev(o["prefixes"]),
                       to_dev(o["child_offsets"]), to_dev(o["parents"]), to_dev(o["level_range"]),
                       to_dev(o["internal_to_leaf"]), to_dev(o["leaf_to_internal"]))
    return v


class GpuImpl:
    which = "cuda"

    # ---- SFC
    def encode(self, kind, ix, iy, iz, kdtype=np.uint64):
        bits = 8 * np.dtype(kdtype).itemsize
        k = ops.encode_sfc(to_dev(np.array([ix], np.uint32)), to_dev(np.array([iy], np.uint32)),
                           to_dev(np.array([iz], np.uint32)), bits, kind)
        return int(keys_np(k, kdtype)[0])

    def decode(self, kind, key, kdtype=np.uint64):
        x, y, z = ops.decode_sfc(to_dev(np.array([key], dtype=kdtype)), kind)
        return (int(x.cpu().numpy().view(np.uint32)[0]), int(y.cpu().numpy().view(np.uint32)[0]),
                int(z.cpu().numpy().view(np.uint32)[0]))

    def sfc_ibox(self, kind, key, level, kdtype=np.uint64):
        """via geo centres in a box whose cells have unit length: centre -/+ size = integer box"""
        ml = 10 if np.dtype(kdtype) == np.uint32 else 21
        box = ops.make_box((0, float(1 << ml)), dtype=torch.float64)
        prefix = np.array([(1 << (3 * level)) | (int(key) >> (3 * (ml - level)))], dtype=kdtype)
        c, s = ops.compute_geo_centers(to_dev(prefix), box, torch.float64, kind)
        c, s = c.cpu().numpy()[0], s.cpu().numpy()[0]
        lo, hi = c - s, c + s
        return (int(lo[0]), int(hi[0]), int(lo[1]), int(hi[1]), int(lo[2]), int(hi[2]))

    def compute_sfc_keys(self, kind, x, y, z, box, kdtype=np.uint64, keys=None):
        bits = 8 * np.dtype(kdtype).itemsize
        dk = None if keys is None else to_dev(keys)
        k = ops.compute_sfc_keys(to_dev(x), to_dev(y), to_dev(z), conv_box(box), dk, bits, kind)
        return keys_np(k, kdtype)

    # ---- sort / gather
    def sort_by_key(self, keys, values=None):
        k = to_dev(keys)
        v = None if values is None else to_dev(values)
        k, o = ops.sort_by_key(k, v)
        return keys_np(k, keys.dtype), o.cpu().numpy().view(np.uint32)

    def gather(self, order, src):
        src = np.ascontiguousarray(src)
        out = ops.gather(to_dev(order), to_dev(src))
        return out.cpu().numpy().view(src.dtype) if src.dtype.kind == "u" else out.cpu().numpy()

    # ---- leaf array
    def compute_node_counts(self, tree, keys, max_count=0xFFFFFFFF, guess=False, counts=None):
        c = None if counts is None else to_dev(counts)
        k = to_dev(keys) if keys.size else torch.empty(0, dtype=to_dev(tree).dtype, device=DEV)
        out = ops.compute_node_counts(to_dev(tree), k, c, max_count, guess)
        return out.cpu().numpy().view(np.uint32)

    def rebalance_decision(self, tree, counts, bucket):
        n = tree.size - 1
        t, c = to_dev(tree), to_dev(counts)
        ops_ = ops.rebalance_decision(t, c, bucket).cpu().numpy()
        scanned, new_n, conv = ops.compute_node_ops(t, c, bucket)
        s = scanned.cpu().numpy()
        assert new_n == int(ops_.sum()) and np.array_equal(np.diff(s), ops_) and conv == bool((ops_ == 1).all())
        return np.concatenate([ops_, [0]]).astype(np.int32), conv

    def rebalance_tree(self, tree, ops_):
        n = tree.size - 1
        scan = np.zeros(n + 1, dtype=np.int32)
        np.cumsum(ops_[:n], out=scan[1:])
        new_n = int(scan[n])
        nt = ops.rebalance_tree(to_dev(tree), to_dev(scan), new_n)
        return keys_np(nt, tree.dtype), scan

    def _leaf_array(self, kd):
        return ops.LeafArray(8 * np.dtype(kd).itemsize, capacity=1024, device=DEV)

    def update_octree(self, keys, bucket, tree, counts, max_count=0xFFFFFFFF):
        la = self._leaf_array(tree.dtype)
        la.set(to_dev(tree), to_dev(counts))
        conv = la.update(to_dev(keys), bucket, max_count)
        return keys_np(la.tree, tree.dtype), la.counts.cpu().numpy().view(np.uint32), conv

    def compute_octree(self, keys, bucket, max_count=0xFFFFFFFF):
        la = self._leaf_array(keys.dtype)
        la.compute(to_dev(keys), bucket, max_count)
        return keys_np(la.tree, keys.dtype), la.counts.cpu().numpy().view(np.uint32)

    # ---- internal tree
    def build_octree(self, leaves):
        v = ops.build_octree(to_dev(leaves))
        torch.cuda.synchronize()
        return dict(prefixes=keys_np(v.prefixes, leaves.dtype), child_offsets=v.child_offsets.cpu().numpy(),
                    parents=v.parents.cpu().numpy(), level_range=v.level_range.cpu().numpy(),
                    internal_to_leaf=v.internal_to_leaf.cpu().numpy(),
                    leaf_to_internal=v.leaf_to_internal.cpu().numpy(), num_leaf_nodes=v.num_leaf_nodes,
                    num_internal_nodes=v.num_internal_nodes, num_nodes=v.num_nodes)

    def upsweep_sum(self, level_range, child_offsets, counts):
        c = to_dev(counts)
        ops.upsweep_sum(level_range, to_dev(child_offsets), c)
        return c.cpu().numpy().view(np.uint32)

    def geo_centers(self, kind, prefixes, box, fdtype=np.float64):
        td = torch.float32 if np.dtype(fdtype) == np.float32 else torch.float64
        c, s = ops.compute_geo_centers(to_dev(prefixes), conv_box(box), td, kind)
        return c.cpu().numpy(), s.cpu().numpy()

    # ---- halos
    def halo_radii(self, h, segments, scale=2.0):
        return ops.halo_radii(to_dev(h), to_dev(segments), scale).cpu().numpy()

    def find_halos(self, kind, octree, leaves, radii, box, first, last, fdtype=np.float64):
        f = ops.find_halos(oct_to_dev(octree), to_dev(leaves), to_dev(radii), conv_box(box), first, last, None, kind)
        return f.cpu().numpy()

    def compute_node_layout(self, counts, flags, first, last):
        return ops.compute_node_layout(to_dev(counts), to_dev(flags), first, last).cpu().numpy().view(np.uint32)

    # ---- neighbours
    def find_neighbors(self, x, y, z, h, first, last, box, octree, layout, centers, sizes, ngmax, ext=1.0):
        nb, nc = ops.find_neighbors(to_dev(x), to_dev(y), to_dev(z), to_dev(h), first, last, conv_box(box),
                                    oct_to_dev(octree), to_dev(layout), to_dev(centers), to_dev(sizes), ngmax,
                                    search_ext_factor=ext)
        return nb.cpu().numpy().view(np.uint32), nc.cpu().numpy().view(np.uint32)
