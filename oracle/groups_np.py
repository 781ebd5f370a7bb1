"""TEST INFRASTRUCTURE ONLY — numpy restatement of the reference's target-group construction.

The reference implements these on the GPU only (no CPU twin, nothing to compile into oracle/_ref):
  computeFixedGroups / fixedGroupsKernel     [ref: traversal/groups_gpu.cu:39-69]
  findSplits                                 [ref: traversal/groups_gpu.cuh:54-95]
  makeSplits                                 [ref: traversal/groups_gpu.cuh:110-137]
  groupSplitsKernel                          [ref: traversal/groups_gpu.cuh:164-232]
  computeGroupSplitsImpl (scan / lengths / scan)  [ref: traversal/groups_gpu.cu:72-116]
Pinned on the known-answer vectors of test/unit_cuda/traversal/groups.cu (tests/test_groups_oracle.py).  Arithmetic is
IEEE without contraction, in the types the reference uses (coordinates Tc, leaf volume in the type of h)."""
import numpy as np

WARP = 32


def fixed_groups(first, last, group_size):
    n = -(-(last - first) // group_size)
    g = first + group_size * np.arange(n + 1, dtype=np.int64)
    g[n] = last
    return g.astype(np.uint32)


def make_splits(masks):
    """masks: iterable of 32-bit words (segment 0 first) -> list of lengths (popcount + 1 entries)"""
    out, carry = [], 0
    for m in masks:
        m = int(m)
        remaining = 32
        while m:
            length = (m & -m).bit_length()  # trailing zeros + 1
            remaining -= length
            out.append(length + carry)
            carry = 0
            m >>= length
        carry += remaining
    out.append(carry)
    return out


def _tree_level(rng, max_level):
    return max_level - (int(rng).bit_length() - 1) // 3


def group_split_masks(first, last, x, y, z, leaves, layout, inv_lengths, group_size, tol_factor, th=np.float64):
    """-> (masks uint32[num_fixed, group_size/32], num_splits uint32[num_fixed])"""
    tc = x.dtype.type
    th = np.dtype(th).type
    nwt = group_size // WARP
    num_fixed = -(-(last - first) // group_size)
    max_level = 10 if leaves.dtype == np.uint32 else 21
    num_leaves = leaves.size - 1
    masks = np.zeros((num_fixed, nwt), dtype=np.uint32)
    lanes = np.arange(group_size, dtype=np.int64)
    tol = np.float32(tol_factor)
    for g in range(num_fixed):
        body = np.minimum(first + g * group_size + lanes, last - 1)
        # leaf edge of the first warp-width of bodies only (the reference indexes leafIdx[0] for every segment)
        leaf = np.searchsorted(layout[:num_leaves], body[:WARP], side="right") - 1
        vol = th(1)
        for l in np.unique(leaf):
            level = _tree_level(int(leaves[l + 1]) - int(leaves[l]), max_level)
            half = th(1 << (max_level - level)) * (th(0.5) / th(1 << max_level))
            vol = min(vol, th(8) * half * half * half)
        dist = tc(np.cbrt(vol) * (th(tol) if th is np.float64 else tol))
        crit2 = tc(dist * dist)
        px = (x[body] * tc(inv_lengths[0])).astype(tc)
        py = (y[body] * tc(inv_lengths[1])).astype(tc)
        pz = (z[body] * tc(inv_lengths[2])).astype(tc)
        nxt = np.minimum(lanes + 1, group_size - 1)  # last lane of the last segment compares with itself
        dx, dy, dz = px[nxt] - px, py[nxt] - py, pz[nxt] - pz
        d2 = (dx * dx + dy * dy) + dz * dz
        bits = d2 > crit2
        for k in range(nwt):
            seg = bits[k * WARP:(k + 1) * WARP]
            masks[g, k] = np.uint32(int(np.sum(seg.astype(np.uint64) << np.arange(WARP, dtype=np.uint64))))
    num_splits = np.array([1 + sum(bin(int(w)).count("1") for w in row) for row in masks], dtype=np.uint32)
    return masks, num_splits


def group_splits(first, last, x, y, z, leaves, layout, inv_lengths, group_size, tol_factor, th=np.float64):
    """computeGroupSplits: group starts (num_groups + 1 entries, last one = `last`)"""
    if last <= first:
        return np.array([last], dtype=np.uint32)
    masks, _ = group_split_masks(first, last, x, y, z, leaves, layout, inv_lengths, group_size, tol_factor, th)
    lengths = []
    for row in masks:
        lengths += make_splits(row)
    starts = first + np.concatenate(([0], np.cumsum(np.array(lengths, dtype=np.int64))))
    starts[-1] = last  # the reference overwrites the last entry after the second scan
    return starts.astype(np.uint32)
