"""TEST INFRASTRUCTURE ONLY — numpy restatement of the reference's small device primitives (integer / byte / element-wise
work), used by tests/ as the checker of cornerstone_octree_b200/csrc/primitives.cu.  Each function cites the reference
lines it follows (paths relative to the reference's include/cstone/).  The reference implements them with thrust / one
trivial kernel each; their CPU meaning is the std:: algorithm of the same name."""
import numpy as np


def fill(n, value, dtype):
    """[ref: primitives/primitives_gpu.cu:51-62 fillGpu = thrust::fill]"""
    return np.full(n, value, dtype=dtype)


def scale(a, value):
    """[ref: primitives/primitives_gpu.cu:64-84 scaleGpu: first[i] *= value, in the element type]"""
    return (a * a.dtype.type(value)).astype(a.dtype)


def increment(a, value):
    """[ref: primitives/primitives_gpu.cu:86-107 incrementGpu: d_first[i] = first[i] + value, wrapping unsigned]"""
    with np.errstate(over="ignore"):
        return (a + a.dtype.type(value)).astype(a.dtype)


def max_norm_square(x, y, z):
    """[ref: primitives/primitives_gpu.cu:194-215] max of x*x + y*y + z*z (left to right, element type), init 0"""
    if x.size == 0:
        return x.dtype.type(0)
    v = (x * x + y * y) + z * z
    return max(x.dtype.type(0), v.max())


def lower_bound(a, values):
    """[ref: primitives/primitives_gpu.cu:217-238 lowerBoundGpu = thrust::lower_bound]"""
    return np.searchsorted(a, values, side="left")


def segment_max(inp, segments, out_dtype):
    """[ref: primitives/primitives_gpu.cu:240-275 segmentMaxKernel: localMax starts at 0, output converted to Tout]"""
    out = np.zeros(segments.size - 1, dtype=out_dtype)
    for s in range(segments.size - 1):
        b, e = int(segments[s]), int(segments[s + 1])
        m = inp.dtype.type(0)
        if e > b:
            m = max(m, inp[b:e].max())
        out[s] = out_dtype(m)
    return out


def reduce_sum(a, init):
    """[ref: primitives/primitives_gpu.cu:277-283 reduceGpu<unsigned, size_t>: 64-bit sum]"""
    return int(init) + int(a.astype(np.uint64).sum())


def count(a, value):
    """[ref: primitives/primitives_gpu.cu:440-448 countGpu = thrust::count]"""
    return int((a == a.dtype.type(value)).sum())


def sort_keys(a):
    """[ref: primitives/primitives_gpu.cu:295-324 sortGpu: ascending]"""
    return np.sort(a, kind="stable")


def gather_ranges(range_scan, range_offsets, src, buffer_size):
    """[ref: halos/gather_halos_gpu.cu:25-42 gatherRangesKernel]: packed index t belongs to the last range whose
    scan value is <= t (upper_bound - 1)"""
    t = np.arange(buffer_size)
    r = np.searchsorted(range_scan, t, side="right") - 1
    return src[range_offsets[r].astype(np.int64) + t - range_scan[r].astype(np.int64)]


def count_fill_sfc_gaps(oracle, tree):
    """[ref: tree/csarray_gpu.cu:231-265 countSfcGapsGpu / fillSfcGapsGpu; tree/csarray.hpp:508-531 computeSpanningTree]
    with spanSfcRange taken from the C oracle (port or reference build)"""
    n = tree.size - 1
    pieces = [oracle.span_sfc_range(tree[i], tree[i + 1], tree.dtype) for i in range(n)]
    ops = np.array([p.size for p in pieces], dtype=np.int32)
    new_tree = np.concatenate(pieces + [tree[-1:]]).astype(tree.dtype)
    return ops, new_tree
