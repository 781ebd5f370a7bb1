"""Known-answer vectors of the reference's own unit tests for the focus-tree functions, written against a tiny
interface so that the reference build (CPU) and the CUDA library run the same checks
[ref: test/unit/focus/octree_focus.cpp:244-310 TEST(FocusedOctree, keyEnforcement), :208-236 nodeOpsKeepAlive].

impl.enforce(keys, octree, node_ops) -> (node_ops, status)     status: 0 converged, 1 cancelMerge, 2 rebalance, 3 failed
impl.protect(octree, node_ops)       -> (node_ops, converged)"""
import numpy as np


def _tree_divide_octant(octant, kd=np.uint32):
    """OctreeMaker{}.divide().divide(octant): root split, one level-1 octant split again"""
    ml = 10 if np.dtype(kd) == np.uint32 else 21
    r1, r2 = 1 << (3 * (ml - 1)), 1 << (3 * (ml - 2))
    keys = set(i * r1 for i in range(9)) | set(octant * r1 + i * r2 for i in range(8))
    return np.array(sorted(keys), dtype=kd), ml


def _key(digits, ml):
    """decodePlaceholderBit(0<digits>): the key whose leading octal digits are `digits`"""
    v = 0
    for d in digits:
        v = v * 8 + d
    return v << (3 * (ml - len(digits)))


def check_key_enforcement(impl, build_octree):
    kd = np.uint32
    tree, ml = _tree_divide_octant(1, kd)
    oc = build_octree(tree)
    assert oc["num_nodes"] == 17
    # | 0 | 1 2 3 4 5 6 7 8 |  node 2 has the children 9..16
    base = np.array([1, 1] + [0] * 15, dtype=np.int32)

    ops, st = impl.enforce(np.array([_key((1, 1), ml)], dtype=kd), oc, base)
    assert st == 1 and ops.tolist() == [1] * 17  # cancelMerge: an existing node, every merge above it undone

    ops, st = impl.enforce(np.array([_key((1, 1, 2), ml)], dtype=kd), oc, base)
    assert st == 2 and ops.tolist() == [1] * 10 + [8] + [1] * 6  # rebalance: node 10 is split

    ops, st = impl.enforce(np.array([_key((0, 1), ml)], dtype=kd), oc, base)
    ops, _ = impl.protect(oc, ops)
    assert st == 2 and ops.tolist() == [1, 8] + [1] * 8 + [0] * 7

    # splitting node 1 does not invalidate the merge of nodes 9-16; a key two levels down cannot be resolved in one step
    start = np.array([1] * 10 + [0] * 7, dtype=np.int32)
    ops, st = impl.enforce(np.array([_key((0, 1), ml)], dtype=kd), oc, start)
    assert st == 2
    ops, st = impl.enforce(np.array([_key((0, 1, 1), ml)], dtype=kd), oc, ops)
    assert st == 3 and ops.tolist() == [1, 8] + [1] * 8 + [0] * 7
    # both keys at once: the status is the maximum
    ops, st = impl.enforce(np.array([_key((0, 1), ml), _key((0, 1, 1), ml)], dtype=kd), oc, start)
    assert st == 3 and ops.tolist() == [1, 8] + [1] * 8 + [0] * 7


def check_protect_ancestors(impl, build_octree):
    """nodeOpsKeepAlive: a non-zero op in a subtree keeps the merge of its ancestors from happening"""
    for kd in (np.uint32, np.uint64):
        ml = 10 if kd == np.uint32 else 21
        r1, r2 = 1 << (3 * (ml - 1)), 1 << (3 * (ml - 2))
        keys = set(i * r1 for i in range(9)) | set(i * r2 for i in range(8)) | set(7 * r1 + i * r2 for i in range(8))
        oc = build_octree(np.array(sorted(keys), dtype=kd))
        assert oc["num_nodes"] == 25
        ops = np.zeros(25, dtype=np.int32)
        ops[0] = 1
        ops[9] = 1  # a kept leaf below node 1
        out, conv = impl.protect(oc, ops)
        want = [0] * 25
        want[0] = want[1] = want[9] = 1
        assert out.tolist() == want and not conv


def check_rebalance_decision(impl, build_octree, upsweep_sum):
    """TEST(FocusedOctree, rebalanceDecision), first two cases [ref: test/unit/focus/octree_focus.cpp:80-130]:
    impl.decision(octree, counts, macs, focus_start, focus_end, bucket) -> node ops; leaf ops after protectAncestors"""
    for kd in (np.uint32, np.uint64):
        ml = 10 if kd == np.uint32 else 21
        r1, r2 = 1 << (3 * (ml - 1)), 1 << (3 * (ml - 2))
        keys = set(i * r1 for i in range(9)) | set(i * r2 for i in range(8)) | set(7 * r1 + i * r2 for i in range(8))
        tree = np.array(sorted(keys), dtype=kd)  # OctreeMaker{}.divide().divide(0).divide(7): 22 leaves
        oc = build_octree(tree)
        L, I = oc["num_leaf_nodes"], oc["num_internal_nodes"]
        assert L == 22
        l2i = oc["leaf_to_internal"][I:]
        cases = [
            # counts win: the leaves 14-21 are fused although the MAC asks for a split of 14
            ([1, 1, 1, 2, 1, 1, 1, 1, 1, 1, 2, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0, 0],
             [0, 0, 1, 0, 1, 1, 1, 1, 1, 0, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0],
             [1, 1, 1, 8, 1, 1, 1, 1, 1, 1, 8, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0], False),
            # the MAC wins: the leaves 14-21 stay, but are not split
            ([1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 2, 1, 0, 0, 0, 0],
             [0, 0, 1, 1, 1, 1, 1, 1, 1, 0, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0, 0],
             [1] * 22, True),
        ]
        for leaf_counts, leaf_macs, want, want_conv in cases:
            counts = np.zeros(oc["num_nodes"], dtype=np.uint32)
            counts[l2i] = leaf_counts
            counts = upsweep_sum(oc["level_range"], oc["child_offsets"], counts)
            macs = np.zeros(oc["num_nodes"], dtype=np.int8)
            macs[l2i] = leaf_macs
            for prefix in (0o1, 0o10, 0o17):  # internal cell macs in placeholder-bit format
                (idx,) = np.nonzero(oc["prefixes"][:oc["num_nodes"]] == prefix)
                macs[idx[0]] = 1
            node_ops = impl.decision(oc, counts, macs, int(tree[0]), int(tree[8]), 1)
            node_ops, conv = impl.protect(oc, node_ops)
            assert node_ops[l2i].tolist() == want
            assert bool(conv) == want_conv
