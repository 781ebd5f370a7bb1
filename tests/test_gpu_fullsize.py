"""BASELINE.json's configurations at their full per-GPU sizes, checked through size-independent properties (the CPU
oracle would take minutes there): sortedness, keys reproducible from the output coordinates, permutation-invariant
integer checksums, fields travelling with their particles, leaf counts re-derived with searchsorted, bucket invariant,
idempotence of a second sync, neighbour lists of sampled targets against an all-pairs search in torch (same double
expression, no FMA contraction across torch's elementwise kernels)."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NG0 = 100


def _need(gb):
    import torch
    free, _ = torch.cuda.mem_get_info()
    if free < gb * 2 ** 30:
        pytest.skip("needs %d GB of free device memory" % gb)


def _uniform(n, seed, dtype):
    import torch
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    x, y, z = (torch.rand(n, generator=g, device="cuda", dtype=dtype) for _ in range(3))
    h = torch.full((n,), 0.5 * (3.0 * NG0 / (4.0 * math.pi * n)) ** (1.0 / 3.0), device="cuda", dtype=dtype)
    return x, y, z, h


def _plummer(n, seed):
    """Plummer sphere drawn on the device (same recipe as tests/gen.py), h from the local density"""
    import torch
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    u = lambda: torch.rand(n, generator=g, device="cuda", dtype=torch.float64)
    a = 3.0 * math.pi / 16.0
    m = u().clamp_(1e-10, 0.999)
    r = a / torch.sqrt(m.pow(-2.0 / 3.0) - 1.0)
    ct = 2.0 * u() - 1.0
    st = torch.sqrt(1.0 - ct * ct)
    ph = 2.0 * math.pi * u()
    x, y, z = r * st * torch.cos(ph), r * st * torch.sin(ph), r * ct
    del m, ct, st, ph
    rho = n * 3.0 / (4.0 * math.pi * a ** 3) * (1.0 + (r / a) ** 2) ** -2.5
    h = 0.5 * (3.0 * NG0 / (4.0 * math.pi * rho)) ** (1.0 / 3.0)
    lim = float(torch.max(r)) * 1.0001
    h = torch.minimum(h, torch.tensor(0.05 * 2 * lim, device="cuda", dtype=torch.float64))
    return x, y, z, h, lim


def _checksum(t):
    """permutation-invariant: wrapping integer sum and xor-free second moment of the bit patterns"""
    import torch
    v = t.view(torch.int64) if t.element_size() == 8 else t.view(torch.int32).to(torch.int64)
    return int(v.sum().item()), int((v * 0x9E3779B1 >> 7).sum().item())


def _check_tree(tarr, keys, bucket, key_bits):
    import torch
    leaves = tarr["leaves"]
    counts = tarr["counts"].to(torch.int64)
    lv = leaves.view(torch.int64) if key_bits == 64 else leaves.view(torch.int32).to(torch.int64)
    kv = keys.view(torch.int64) if key_bits == 64 else keys.view(torch.int32).to(torch.int64)
    max_level = 21 if key_bits == 64 else 10
    end = 1 << (3 * max_level)
    assert int(lv[0]) == 0 and int(lv[-1]) & 0xFFFFFFFFFFFFFFFF == end  # 2^63 reads as int64 min
    width = lv[1:] - lv[:-1]  # wraps to the right positive value for the last 64-bit leaf
    lv = lv.clone()
    lv[-1] = min(end, 2 ** 63 - 1)
    assert bool((lv[1:] > lv[:-1]).all())
    pos = torch.searchsorted(kv, lv)
    pos[-1] = keys.numel()
    recount = pos[1:] - pos[:-1]
    assert bool((recount == counts).all()) and int(counts.sum()) == keys.numel()
    # a leaf above the bucket size is legal only at the maximum depth
    assert bool(((counts <= bucket) | (width == 1)).all())
    # every width is a power of 8 and the start is aligned to it (cornerstone invariant)
    lg = torch.log2(width.to(torch.float64)).round().to(torch.int64)
    assert bool(((lg % 3 == 0) & (torch.bitwise_left_shift(torch.ones_like(lg), lg) == width) & (lv[:-1] % width == 0)).all())


def _check_neighbors(x, y, z, h, first, last, nb, nc, ngmax, samples, seed, box_len=None):
    """sampled targets against all pairs; positions are the arrays findNeighbors searched"""
    import torch
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    idx = torch.randint(first, last, (samples,), generator=g).tolist()
    for i in idx:
        dx, dy, dz = x - x[i], y - y[i], z - z[i]
        if box_len is not None:
            dx -= box_len * torch.round(dx / box_len)
            dy -= box_len * torch.round(dy / box_len)
            dz -= box_len * torch.round(dz / box_len)
        d2 = dx * dx + dy * dy + dz * dz
        hit = d2 < 4.0 * h[i] * h[i]
        hit[i] = False
        ref = torch.nonzero(hit).flatten()
        c = int(nc[i - first])
        assert c == ref.numel(), (i, c, ref.numel())
        if c <= ngmax:
            got = nb[i - first, :c].to(torch.int64).sort().values
            assert bool((got == ref).all()), i


def _sync_and_check(dom, x, y, z, h, props, key_bits, bucket_focus, ngmax, samples):
    import torch
    from cornerstone_octree_b200 import ops
    n = x.numel()
    sums = [_checksum(t) for t in (x, y, z, h)]
    keys, sx, sy, sz, sh, sp = dom.sync(x, y, z, h, props)
    assert dom.start_index == 0 and dom.end_index == n
    kv = keys.view(torch.int64) if key_bits == 64 else keys.view(torch.int32)
    assert bool((kv[1:] >= kv[:-1]).all())
    assert [_checksum(t) for t in (sx, sy, sz, sh)] == sums
    rekey = ops.compute_sfc_keys(sx, sy, sz, dom.box(), key_bits_=key_bits)
    assert bool((rekey.view(kv.dtype) == kv).all())
    del rekey
    _check_tree(dom.tree_arrays(), keys, bucket_focus, key_bits)
    return keys, sx, sy, sz, sh, sp


def test_c1_32M_uniform_hilbert64():
    """configs[1]: 32 M uniform particles, 64-bit Hilbert keys, bucketSize 64, + findNeighbors ngmax 150"""
    import torch
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200.domain import Domain
    _need(40)
    n = 32_000_000
    x, y, z, h = _uniform(n, 1, torch.float64)
    dom = Domain(0, 1, 65536, 64, 0.5, ops.make_box((0.0, 1.0)))
    keys, sx, sy, sz, sh, _ = _sync_and_check(dom, x, y, z, h, [], 64, 64, 150, 40)
    del x, y, z, h
    nb, nc = dom.find_neighbors(sx, sy, sz, sh, 150)
    assert 95 < float(nc.to(torch.float64).mean()) < 105
    _check_neighbors(sx, sy, sz, sh, 0, n, nb, nc, 150, 40, 11)
    # idempotence: a second sync of the sorted arrays changes nothing
    k2, x2, y2, z2, h2, _ = dom.sync(sx, sy, sz, sh)
    assert bool((k2 == keys).all()) and bool((x2 == sx).all()) and bool((z2 == sz).all())
    dom.close()


def test_c2_64M_plummer_deep_tree():
    """configs[2]: 64 M Plummer-clustered particles, findNeighbors ngmax 150 on a deep unbalanced tree"""
    import torch
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200.domain import Domain
    _need(70)
    n = 64_000_000
    x, y, z, h, lim = _plummer(n, 2)
    dom = Domain(0, 1, 65536, 64, 0.5, ops.make_box((-lim, lim)))
    keys, sx, sy, sz, sh, _ = _sync_and_check(dom, x, y, z, h, [], 64, 64, 150, 24)
    del x, y, z, h
    t = dom.tree_arrays()
    lv = t["leaves"].view(torch.int64)
    depth = 21 - (torch.log2((lv[1:] - lv[:-1]).to(torch.float64).abs()).round().to(torch.int64) // 3)
    assert int(depth.max()) - int(depth.min()) >= 6  # deep and unbalanced
    del t, lv, depth
    nb, nc = dom.find_neighbors(sx, sy, sz, sh, 150)
    _check_neighbors(sx, sy, sz, sh, 0, n, nb, nc, 150, 24, 12)
    dom.close()


def test_c3_128M_with_4_fields():
    """configs[3], one rank's share: 128 M particles (double x,y,z,h + 4 fields) through Domain::sync"""
    import torch
    from cornerstone_octree_b200 import ops
    from cornerstone_octree_b200.domain import Domain
    _need(110)
    n = 128_000_000
    x, y, z, h = _uniform(n, 3, torch.float64)
    fields = [x * 2.0 + 1.0, y - z, torch.arange(n, device="cuda", dtype=torch.int64), (x * 1000.0).to(torch.float32)]
    fsum = [_checksum(f) for f in fields]
    dom = Domain(0, 1, 65536, 64, 0.5, ops.make_box((0.0, 1.0)))
    keys, sx, sy, sz, sh, sp = _sync_and_check(dom, x, y, z, h, fields, 64, 64, 0, 0)
    del x, y, z, h, fields
    assert [_checksum(f) for f in sp] == fsum
    # fields travelled with their particles: they are functions of the coordinates
    assert bool((sp[0] == sx * 2.0 + 1.0).all()) and bool((sp[1] == sy - sz).all())
    assert bool((sp[3] == (sx * 1000.0).to(torch.float32)).all())
    ids = sp[2]
    assert int(ids.min()) == 0 and int(ids.max()) == n - 1
    dom.close()


def test_c4_morton32_bucket16_sort_bound():
    """configs[4], one rank's share (256 M / 8): 32-bit Morton keys, bucketSize 16: 10-level tree, sort-bound"""
    import torch
    from cornerstone_octree_b200 import ops
    _need(20)
    n = 32_000_000
    x, y, z, _ = _uniform(n, 4, torch.float32)
    box = ops.make_box((0.0, 1.0), dtype=torch.float32)
    keys = ops.compute_sfc_keys(x, y, z, box, key_bits_=32, kind=ops.MORTON)
    ksum = _checksum(keys)
    skeys, order = ops.sort_by_key(keys.clone())  # sorts in place
    kv = skeys.view(torch.int32)
    assert bool((kv[1:] >= kv[:-1]).all()) and _checksum(skeys) == ksum
    o64 = order.view(torch.int32).to(torch.int64)
    assert bool((keys.view(torch.int32)[o64] == kv).all())
    # stable: equal keys keep their input order
    same = kv[1:] == kv[:-1]
    assert bool((o64[1:][same] > o64[:-1][same]).all())
    sx, sy, sz = ops.gather_multi(order, [x, y, z])
    assert bool((ops.compute_sfc_keys(sx, sy, sz, box, key_bits_=32, kind=ops.MORTON).view(torch.int32) == kv).all())
    # decode(encode) round trip of the integer coordinates
    ix, iy, iz = ops.decode_sfc(skeys[:1_000_000], kind=ops.MORTON)
    assert bool((ops.encode_sfc(ix, iy, iz, key_bits_=32, kind=ops.MORTON).view(torch.int32) == kv[:1_000_000]).all())
    leaves = ops.LeafArray(32)
    leaves.compute(skeys, 16)
    _check_tree(dict(leaves=leaves.tree, counts=leaves.counts), skeys, 16, 32)
    lv = leaves.tree.view(torch.int32).to(torch.int64)
    assert int((lv[1:] - lv[:-1]).min()) <= 64  # level 8 of the 10 levels a 32-bit key allows
