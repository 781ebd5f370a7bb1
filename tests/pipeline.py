"""Shared test helper: run the whole hot path through ONE implementation with a common interface and
return every intermediate array so that two implementations can be compared stage by stage.

`impl` is an object with the oracle.loader.Oracle method set (the CPU checkers) or the
cornerstone_octree_b200 numpy-facing adapter used by the GPU tests (tests/gpu_adapter.py)."""
import numpy as np


def run(impl, kind, x, y, z, h, box, kdtype, bucket, ngmax, halo_range=None, nb_range=None, ext=1.0):
    out = {}
    fdtype = x.dtype
    keys = impl.compute_sfc_keys(kind, x, y, z, box, kdtype)
    out["keys_unsorted"] = keys
    skeys, order = impl.sort_by_key(keys)
    out["keys"], out["order"] = skeys, order
    xs, ys, zs, hs = (impl.gather(order, a) for a in (x, y, z, h))
    out["x"], out["y"], out["z"], out["h"] = xs, ys, zs, hs

    tree, counts = impl.compute_octree(skeys, bucket)
    out["tree"], out["counts"] = tree, counts
    oct_ = impl.build_octree(tree)
    for k in ("prefixes", "child_offsets", "parents", "level_range", "internal_to_leaf", "leaf_to_internal"):
        out[k] = oct_[k][:oct_["num_nodes"]] if k in ("child_offsets",) else oct_[k]
    centers, sizes = impl.geo_centers(kind, oct_["prefixes"], box, fdtype)
    out["centers"], out["sizes"] = centers, sizes

    L = tree.size - 1
    first, last = halo_range if halo_range is not None else (L // 4, L // 2)
    seg = np.zeros(L + 1, dtype=np.uint32)
    np.cumsum(counts, out=seg[1:])
    radii = impl.halo_radii(hs, seg, 2.0 * ext)
    # only local leaves carry a radius (halos/halos.hpp:156-158: segmentMax over [firstNode,lastNode))
    mask = np.zeros(L, dtype=bool)
    mask[first:last] = True
    radii = np.where(mask, radii, np.float32(0)).astype(np.float32)
    out["radii"] = radii
    flags = impl.find_halos(kind, oct_, tree, radii, box, first, last, fdtype)
    out["flags"] = flags
    out["layout_halo"] = impl.compute_node_layout(counts, flags, first, last)

    # neighbour search over the full particle set with the plain count layout
    layout = seg
    f, l = nb_range if nb_range is not None else (0, x.size)
    nb, nc = impl.find_neighbors(xs, ys, zs, hs, f, l, box, oct_, layout, centers, sizes, ngmax, ext)
    out["neighbors"], out["neighbors_count"] = nb, nc
    return out


def compare(a, b, ngmax, skip=()):
    """bit-exact comparison of two pipeline results; returns list of mismatching stage names"""
    bad = []
    for k in a:
        if k in skip or k == "neighbors":
            continue
        if a[k].dtype.kind == "f":
            same = a[k].shape == b[k].shape and np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8))
        else:
            same = a[k].shape == b[k].shape and np.array_equal(a[k], b[k])
        if not same:
            bad.append(k)
    if "neighbors" in a and "neighbors" not in skip and "neighbors_count" not in bad:
        nc = np.minimum(a["neighbors_count"], ngmax)
        cols = np.arange(ngmax)[None, :]
        valid = cols < nc[:, None]
        if not np.array_equal(np.where(valid, a["neighbors"], 0), np.where(valid, b["neighbors"], 0)):
            bad.append("neighbors")
    return bad
