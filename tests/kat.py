"""Known-answer tests transcribed from the golden vectors of the reference's own unit tests
(/root/reference/test/unit/...; file:line cited per check).  Every function takes an implementation
object (oracle port, oracle ref, or the CUDA adapter) so that the SAME vectors pin all three."""
import numpy as np

from oracle import loader

MORTON, HILBERT = 0, 1


def pad(prefix, length, kd):
    """[ref: sfc/common.hpp:107-111 pad] prefix of `length` bits padded with zeros to 30 / 63 bits"""
    return np.dtype(kd).type(int(prefix) << (3 * loader.max_level(kd) - length))


def node_range(level, kd):
    return 1 << (3 * (loader.max_level(kd) - level))


class OctreeMaker:
    """[ref: tree/cs_util.hpp:152-213 OctreeMaker] divide(i0, i1, ...) adds the 7 upper siblings below the
    node addressed by the octal path (i0, i1, ...)"""

    def __init__(self, kd):
        self.kd = np.dtype(kd)
        self.ml = loader.max_level(kd)
        self.keys = [0, 1 << (3 * self.ml)]

    def divide(self, *path):
        base = 0
        for lvl, digit in enumerate(path):
            base += digit << (3 * (self.ml - 1 - lvl))
        assert base in self.keys
        for s in range(1, 8):
            k = base + (s << (3 * (self.ml - 1 - len(path))))
            assert k not in self.keys
            self.keys.append(k)
        return self

    def make(self):
        return np.array(sorted(self.keys), dtype=self.kd)


def uniform_tree(level, kd):
    """[ref: tree/cs_util.hpp:93-120 makeUniformNLevelTree]"""
    n = 8 ** level
    r = node_range(level, kd)
    return np.array([i * r for i in range(n)] + [node_range(0, kd)], dtype=kd)


# --------------------------------------------------------------------------- SFC
def check_morton(impl):
    for kd in (np.uint32, np.uint64):
        ml = loader.max_level(kd)
        # test/unit/sfc/morton.cpp:42-50  iMorton(5,3,6) at tree level 3
        sh = ml - 3
        assert impl.encode(MORTON, 5 << sh, 3 << sh, 6 << sh, kd) == pad(0b101011110, 9, kd)
    # morton.cpp:58-69  decode of 340 (32 bit)
    assert impl.decode(MORTON, 340, np.uint32) == (5, 2, 4)
    # morton.cpp:71-89  64-bit decodes
    m = (1 << 21) - 1
    assert impl.decode(MORTON, 0x7FFFFFFFFFFFFFFF, np.uint64) == (m, m, m)
    assert impl.decode(MORTON, 0x1249249241249249, np.uint64)[2] == (1 << 21) - 512 - 1
    assert impl.decode(MORTON, 0b0111 << 60, np.uint64) == (1 << 20, 1 << 20, 1 << 20)
    assert impl.decode(MORTON, 0b0011 << 60, np.uint64) == (0, 1 << 20, 1 << 20)


def check_hilbert(impl):
    h2m = [0, 1, 3, 2, 6, 7, 5, 4]
    for kd in (np.uint32, np.uint64):
        ml = loader.max_level(kd)
        half = (1 << ml) // 2
        # test/unit/sfc/hilbert.cpp:100-128 first-order curve, both corners of each octant
        for xi in range(2):
            for yi in range(2):
                for zi in range(2):
                    for off in (0, half - 1):
                        key = impl.encode(HILBERT, half * xi + off, half * yi + off, half * zi + off, kd)
                        assert h2m[(int(key) >> (3 * (ml - 1))) & 7] == 4 * xi + 2 * yi + zi
        # hilbert.cpp:138-161 continuity: consecutive keys across octant borders are 1 apart in space
        for level in range(1, ml):
            for octant in range(8 if level > 1 else 7):
                last = (octant + 1) * node_range(level, kd) - 1
                a = impl.decode(HILBERT, last, kd)
                b = impl.decode(HILBERT, last + 1, kd)
                assert sum(abs(int(p) - int(q)) for p, q in zip(a, b)) == 1
        # hilbert.cpp:171-198 encode/decode inversion on 1000 random points
        rng = np.random.default_rng(0)
        pts = rng.integers(0, 1 << ml, size=(1000, 3))
        for p in pts:
            key = impl.encode(HILBERT, int(p[0]), int(p[1]), int(p[2]), kd)
            assert impl.decode(HILBERT, key, kd) == tuple(int(v) for v in p)
        # hilbert.cpp:216-264 hilbertIBox: the box of a level-l node contains the decoded start key, is aligned
        for level in range(0, ml + 1, 3):
            cube = (1 << ml) >> level
            for p in pts[:50]:
                key = impl.encode(HILBERT, int(p[0]), int(p[1]), int(p[2]), kd)
                start = (int(key) // node_range(level, kd)) * node_range(level, kd)
                b = impl.sfc_ibox(HILBERT, start, level, kd)
                for d in range(3):
                    assert b[2 * d] == (int(p[d]) // cube) * cube and b[2 * d + 1] == b[2 * d] + cube
        # hilbert.cpp:273-308 sfcNeighbor expectations expressed through encode: lower corners of the L1 octants
        # (sfcNeighbor = enclosingBoxCode(key of the shifted lower corner, level))
        top = lambda key: int(key) >> (3 * (ml - 1))
        assert top(impl.encode(HILBERT, 0, half, half, kd)) == 0o2
        assert top(impl.encode(HILBERT, 0, half, 0, kd)) == 0o3
        assert top(impl.encode(HILBERT, half, half, half, kd)) == 0o5


def check_sfc_keys_box(impl):
    """test/unit/sfc/sfc.cpp:37-75 — keys from coordinates equal keys from the integer grid position"""
    for kd in (np.uint32, np.uint64):
        for fd in (np.float32, np.float64):
            ml = loader.max_level(kd)
            box = loader.make_box((-1, 1, 0, 2, 10, 12), dtype=fd)
            n = 200
            rng = np.random.default_rng(3)
            ijk = rng.integers(0, min(1 << ml, 1 << 20), size=(n, 3))
            scale = fd(2.0) / fd(1 << ml)
            x = (fd(-1) + (ijk[:, 0] + fd(0.5)) * scale).astype(fd)
            y = (fd(0) + (ijk[:, 1] + fd(0.5)) * scale).astype(fd)
            z = (fd(10) + (ijk[:, 2] + fd(0.5)) * scale).astype(fd)
            if fd == np.float32 and kd == np.uint64:
                continue  # cell width below float32 resolution at offset 10
            for kind in (MORTON, HILBERT):
                keys = impl.compute_sfc_keys(kind, x, y, z, box, kd)
                want = [impl.encode(kind, int(a), int(b), int(c), kd) for a, b, c in ijk]
                assert np.array_equal(keys, np.array(want, dtype=kd))


# --------------------------------------------------------------------------- sort / gather
def check_sort(impl):
    """test/unit/primitives/gather.cpp:40-75 — sort_by_key returns sorted keys and the sorting permutation; stable"""
    for kd in (np.uint32, np.uint64):
        keys = np.array([0, 50, 10, 60, 20, 70, 30, 80, 40, 90], dtype=kd)
        sk, order = impl.sort_by_key(keys)
        assert list(order) == [0, 2, 4, 6, 8, 1, 3, 5, 7, 9]
        assert list(sk) == [0, 10, 20, 30, 40, 50, 60, 70, 80, 90]
        vals = np.array([-1, -2, -3, -4, -5, -6, -7, -8, -9, -10], dtype=np.float64)
        assert list(impl.gather(order, vals)) == [-1, -3, -5, -7, -9, -2, -4, -6, -8, -10]
        # stability with heavy duplication
        rng = np.random.default_rng(5)
        keys = rng.integers(0, 17, 5000).astype(kd)
        sk, order = impl.sort_by_key(keys)
        ref = np.argsort(keys, kind="stable").astype(np.uint32)
        assert np.array_equal(order, ref) and np.array_equal(sk, keys[ref])


# --------------------------------------------------------------------------- leaf array
def check_node_counts(impl):
    for kd in (np.uint32, np.uint64):
        # test/unit/tree/csarray.cpp:57-77
        tree = OctreeMaker(kd).divide().divide(0).make()
        t = [int(v) for v in tree]
        codes = np.array([t[1], t[1], t[1] + 10, t[1] + 100, t[2] - 1, t[2] + 1, t[11], t[11] + 2, t[12],
                          t[12] + 1000, t[12] + 2000, t[13] - 10, t[13], t[13] + 1], dtype=kd)
        want = [0, 5, 1, 0, 0, 0, 0, 0, 0, 0, 0, 2, 4, 2, 0]
        assert list(impl.compute_node_counts(tree, codes)) == want
        assert list(impl.compute_node_counts(tree, codes, guess=True,
                                             counts=np.array(want, dtype=np.uint32))) == want
        # maxCount clamp [ref: tree/csarray.hpp:100-102]
        assert list(impl.compute_node_counts(tree, codes, max_count=3)) == [min(w, 3) for w in want]
        # no particles at all
        assert list(impl.compute_node_counts(tree, np.zeros(0, dtype=kd))) == [0] * 15
        # csarray.cpp:83-107: 2 particles each in the very first and very last max-depth node
        nr = node_range(0, kd)
        tree2 = _spanning_like(kd)
        pk = np.array([0, 0, nr - 1, nr - 1], dtype=kd)
        c = impl.compute_node_counts(tree2, pk)
        assert c[0] == 2 and c[-1] == 2 and c.sum() == 4


def check_span_sfc_range(impl):
    """[ref: test/unit/sfc/common.cpp:376-419 spanSfcRange32 / spanSfcRange64]"""
    for kd in (np.uint32, np.uint64):
        span = lambda a, b: [int(v) for v in impl.span_sfc_range(a, b, kd)]
        assert span(0, 0o7777) == [0, 0o1000, 0o2000, 0o3000, 0o4000, 0o5000, 0o6000, 0o7000, 0o7100, 0o7200, 0o7300,
                                   0o7400, 0o7500, 0o7600, 0o7700, 0o7710, 0o7720, 0o7730, 0o7740, 0o7750, 0o7760,
                                   0o7770, 0o7771, 0o7772, 0o7773, 0o7774, 0o7775, 0o7776]
        assert span(1, 0o742) == [1, 2, 3, 4, 5, 6, 7, 0o10, 0o20, 0o30, 0o40, 0o50, 0o60, 0o70, 0o100, 0o200, 0o300,
                                  0o400, 0o500, 0o600, 0o700, 0o710, 0o720, 0o730, 0o740, 0o741]
        assert span(0o41305, 0o46000) == [0o41305, 0o41306, 0o41307, 0o41310, 0o41320, 0o41330, 0o41340, 0o41350,
                                          0o41360, 0o41370, 0o41400, 0o41500, 0o41600, 0o41700, 0o42000, 0o43000,
                                          0o44000, 0o45000]
        assert len(span(0o40000, 0o50000)) == 1 and len(span(0o40000, 0o60000)) == 2
        assert span(int(pad(1, 3, kd)), node_range(0, kd)) == [int(pad(k, 3, kd)) for k in range(1, 8)]
        assert span(0, 0) == [] and span(1, 1) == []


def _spanning_like(kd):
    """cornerstone tree resolving key 1 and key 2^(3L)-1 at maximum depth (same tree computeSpanningTree
    {0,1,R-1,R} yields)"""
    ml = loader.max_level(kd)
    m = OctreeMaker(kd)
    for lvl in range(ml):
        m.divide(*([0] * lvl))
    for lvl in range(1, ml):
        m.divide(*([7] * lvl))
    return m.make()


def check_rebalance(impl):
    for kd in (np.uint32, np.uint64):
        tree = OctreeMaker(kd).divide().divide(0).make()
        # test/unit/tree/csarray.cpp:109-123
        counts = np.array([1, 1, 1, 0, 0, 0, 0, 0, 2, 3, 4, 5, 6, 7, 8], dtype=np.uint32)
        ops, conv = impl.rebalance_decision(tree, counts, 4)
        assert list(ops[:15]) == [1, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 8, 8, 8, 8] and not conv
        # csarray.cpp:132-149 single root
        root = OctreeMaker(kd).make()
        ops, conv = impl.rebalance_decision(root, np.array([1], dtype=np.uint32), 4)
        assert list(ops[:1]) == [1] and conv
        # csarray.cpp:159-187 exhausted resolution stays invariant
        m = OctreeMaker(kd)
        for lvl in range(loader.max_level(kd)):
            m.divide(*([0] * lvl))
        deep = m.make()
        cnt = np.ones(deep.size - 1, dtype=np.uint32)
        cnt[0] = 2
        ops, conv = impl.rebalance_decision(deep, cnt, 1)
        assert (ops[:deep.size - 1] == 1).all() and conv
        # csarray.cpp:190-209 rebalanceTree
        ops = np.array([1, 0, 0, 0, 0, 0, 0, 0, 1, 8, 1, 1, 1, 1, 8, 0], dtype=np.int32)
        new_tree, _ = impl.rebalance_tree(tree, ops)
        assert np.array_equal(new_tree, OctreeMaker(kd).divide().divide(2).divide(7).make())
        # multi-level split ops [ref: tree/csarray.hpp:303-306]
        cnt = np.array([4 * 512 + 1], dtype=np.uint32)
        ops, _ = impl.rebalance_decision(root, cnt, 4)
        assert ops[0] == 4096
        cnt[0] = 4 * 64 + 1
        assert impl.rebalance_decision(root, cnt, 4)[0][0] == 512
        cnt[0] = 4 * 8 + 1
        assert impl.rebalance_decision(root, cnt, 4)[0][0] == 64
        new_tree, _ = impl.rebalance_tree(root, np.array([64, 0], dtype=np.int32))
        assert np.array_equal(new_tree, uniform_tree(2, kd))


def check_compute_octree(impl):
    """test/unit/tree/csarray.cpp:304-349 — converged tree on Gaussian keys satisfies the invariants:
    recount matches, every leaf <= bucket (unless at max depth), no mergeable sibling group"""
    import gen
    for kd in (np.uint32, np.uint64):
        keys = gen.gaussian_keys(100000, kd)
        for bucket in (64, 1024, 10000):
            tree, counts = impl.compute_octree(keys, bucket)
            assert tree[0] == 0 and int(tree[-1]) == node_range(0, kd)
            d = np.diff(tree.astype(object))
            assert all(int(v) > 0 and (int(v) & (int(v) - 1)) == 0 and (int(v).bit_length() - 1) % 3 == 0 for v in d)
            lb = np.searchsorted(keys, tree)
            assert np.array_equal(np.diff(lb).astype(np.uint32), counts)
            assert counts.sum() == keys.size
            assert (counts <= bucket).all()
            ops, conv = impl.rebalance_decision(tree, counts, bucket)
            assert conv


# --------------------------------------------------------------------------- internal tree
def check_octree(impl):
    for kd in (np.uint32, np.uint64):
        ml = loader.max_level(kd)
        # test/unit/tree/octree.cpp:92-110 root only
        o = impl.build_octree(OctreeMaker(kd).make())
        assert o["num_nodes"] == 1 and o["prefixes"][0] == 1 and o["child_offsets"][0] == 0
        assert list(o["level_range"][:3]) == [0, 1, 1]
        # octree.cpp irregular tree L3: divide().divide(0).divide(0,2).divide(3)
        tree = OctreeMaker(kd).divide().divide(0).divide(0, 2).divide(3).make()
        o = impl.build_octree(tree)
        L, I = tree.size - 1, (tree.size - 2) // 7
        assert (L, I) == (29, 4) and o["num_nodes"] == 33
        # level-major ordering: prefixes sorted, ranges 1 / 8 / 16 / 8 nodes
        assert list(o["level_range"][:5]) == [0, 1, 9, 25, 33]
        p = o["prefixes"].astype(object)
        assert all(p[i] < p[i + 1] for i in range(32))
        co = o["child_offsets"]
        # root -> level-1 block; node(0) and node(3) internal at level 1; node(0,2) internal at level 2
        assert co[0] == 1 and co[1] == 9 and co[4] == 17 and co[11] == 25
        assert sorted(np.nonzero(co[:33])[0].tolist()) == [0, 1, 4, 11]
        assert list(o["parents"][:4]) == [0, 1, 4, 11]
        # leaf <-> internal maps are inverse and leaf keys line up
        i2l, l2i = o["internal_to_leaf"], o["leaf_to_internal"]
        for node in range(33):
            if co[node] == 0:
                leaf = i2l[node]
                assert 0 <= leaf < L and l2i[leaf + I] == node
                lvl = (int(p[node]).bit_length() - 1) // 3
                assert (int(p[node]) ^ (1 << (3 * lvl))) << (3 * (ml - lvl)) == int(tree[leaf])
            else:
                assert i2l[node] < 0
        # octree.cpp:389-393 upsweep sum on a regular 4x4x4 tree
        o = impl.build_octree(uniform_tree(2, kd))
        cnt = np.zeros(o["num_nodes"], dtype=np.uint32)
        cnt[9:] = 1
        s = impl.upsweep_sum(o["level_range"], o["child_offsets"], cnt)
        assert s[0] == 64 and (s[1:9] == 8).all()
        big = np.zeros(o["num_nodes"], dtype=np.uint32)
        big[9:] = 0xFFFFFFFF
        assert impl.upsweep_sum(o["level_range"], o["child_offsets"], big)[0] == 0xFFFFFFFF


# --------------------------------------------------------------------------- halos
def check_find_halos(impl):
    for kd in (np.uint32, np.uint64):
        # test/unit/traversal/discovery.cpp:62-103 — 4x4x4 grid, radius 0.1 < cell 0.25: the lower half of the SFC
        # touches exactly 16 leaves of the upper half and vice versa
        tree = uniform_tree(2, kd)
        box = loader.make_box((0, 1))
        o = impl.build_octree(tree)
        radii = np.full(64, 0.1, dtype=np.float32)
        f = impl.find_halos(HILBERT, o, tree, radii, box, 0, 32)
        assert f.sum() == 16 and f[:32].sum() == 0
        assert np.array_equal(f, _halos_all2all(impl, tree, radii, box, 0, 32, kd))
        f = impl.find_halos(HILBERT, o, tree, radii, box, 32, 64)
        assert f.sum() == 16 and f[32:].sum() == 0
        assert np.array_equal(f, _halos_all2all(impl, tree, radii, box, 32, 64, kd))
        # periodic box: the halo wraps around
        pbox = loader.make_box((0, 1), (1, 1, 1))
        f = impl.find_halos(HILBERT, o, tree, radii, pbox, 0, 32)
        assert np.array_equal(f, _halos_all2all(impl, tree, radii, pbox, 0, 32, kd))
        assert f.sum() > 16
        # everything local: nothing to flag
        assert impl.find_halos(HILBERT, o, tree, radii, box, 0, 64).sum() == 0


def _halos_all2all(impl, tree, radii, box, first, last, kd):
    """[ref: test/unit/traversal/collisions_a2a.hpp:50-78] brute force box overlap of every local leaf's halo box
    against every non-local leaf"""
    ml = loader.max_level(kd)
    R = 1 << ml
    L = tree.size - 1
    boxes = []
    for i in range(L):
        rng_ = int(tree[i + 1]) - int(tree[i])
        level = ml - (rng_.bit_length() - 1) // 3
        boxes.append(impl.sfc_ibox(HILBERT, int(tree[i]), level, kd))
    flags = np.zeros(L, dtype=np.int32)

    def ov1(a, b, c, d):
        return b > c and d > a

    def ov(a, b, c, d):
        return ov1(a, b, c, d) or ov1(a + R, b + R, c, d) or ov1(a, b, c + R, d + R)

    for i in range(first, last):
        hb = []
        for d in range(3):
            cells = min(int(np.ceil(np.float64(radii[i]) * box.inv_lengths[d] * R)), R - 1)
            lo, hi = boxes[i][2 * d] - cells, boxes[i][2 * d + 1] + cells
            if box.boundaries[d] != 1:
                lo, hi = min(max(0, lo), R), min(max(0, hi), R)
            hb += [lo, hi]
        for j in list(range(0, first)) + list(range(last, L)):
            b = boxes[j]
            if all(ov(b[2 * d], b[2 * d + 1], hb[2 * d], hb[2 * d + 1]) for d in range(3)):
                flags[j] = 1
    return flags


# --------------------------------------------------------------------------- neighbours
def check_find_neighbors(impl, n=2500, fd=np.float64):
    """test/unit/neighbors/findneighbors.cpp:59-151 — tree search equals the O(N^2) search after sorting the
    lists; ngmax = n; uniform and Gaussian clouds, two radii, two boxes, open and periodic"""
    import gen
    import pipeline
    for kd in (np.uint32, np.uint64):
        for limits in ((0., 1., 0., 1., 0., 1.), (-1.2, 0.23, -0.213, 3.213, -5.1, 1.23)):
            for pbc in (0, 1):
                for cloud in ("uniform", "gaussian"):
                    for radius in (0.124, 0.0624):
                        box = loader.make_box(limits, (pbc,) * 3, fd)
                        x, y, z = (gen.uniform if cloud == "uniform" else gen.gaussian)(n, limits, fd)
                        h = np.full(n, radius / 2, dtype=fd)
                        r = pipeline.run(impl, HILBERT, x, y, z, h, box, kd, 64, n, halo_range=(0, 1))
                        _compare_all2all(r, box, n)


def _compare_all2all(r, box, ngmax):
    x, y, z, h = r["x"], r["y"], r["z"], r["h"]
    n = x.size
    L = np.array([box.lengths[d] for d in range(3)])
    pbc = np.array([box.boundaries[d] == 1 for d in range(3)])
    P = np.stack([x, y, z], axis=1).astype(np.float64)
    for i in range(0, n, 1):
        d = P - P[i]
        d -= np.where(pbc, L * np.rint(d / L), 0.0)
        d2 = (d * d).sum(axis=1)
        r2 = (2 * float(h[i])) ** 2
        # skip pairs within rounding distance of the cut-off (the O(N^2) check runs in float64 numpy)
        band = np.abs(d2 - r2) < 1e-9 * r2 if x.dtype == np.float64 else np.abs(d2 - r2) < 1e-5 * r2
        want = set(np.nonzero((d2 < r2) & ~band)[0].tolist()) - {i}
        maybe = set(np.nonzero(band)[0].tolist())
        got = set(r["neighbors"][i, :r["neighbors_count"][i]].tolist())
        assert want <= got and got <= (want | maybe), (i, sorted(want ^ got))
