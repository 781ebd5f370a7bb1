"""Seeded synthetic particle clouds (semantics of the reference's test/coord_samples generators:
uniform in the box, Gaussian clamped to the box [random.hpp:157-174], Plummer sphere [plummer.hpp:16-79])."""
import numpy as np


def uniform(n, limits=(0.0, 1.0), dtype=np.float64, seed=42):
    rng = np.random.default_rng(seed)
    lim = np.asarray(limits, dtype=np.float64)
    if lim.size == 2:
        lim = np.tile(lim, 3)
    out = []
    for d in range(3):  # all x, then all y, then all z from one generator
        out.append((lim[2 * d] + (lim[2 * d + 1] - lim[2 * d]) * rng.random(n)).astype(dtype))
    for d in range(3):  # guard against rounding up to the upper limit in float32
        np.clip(out[d], dtype(lim[2 * d]), np.nextafter(dtype(lim[2 * d + 1]), dtype(lim[2 * d])), out=out[d])
    return out


def gaussian(n, limits=(-1.0, 1.0), dtype=np.float64, seed=42):
    rng = np.random.default_rng(seed)
    lim = np.asarray(limits, dtype=np.float64)
    if lim.size == 2:
        lim = np.tile(lim, 3)
    out = []
    for d in range(3):
        lo, hi = lim[2 * d], lim[2 * d + 1]
        v = rng.normal((hi + lo) / 2, (hi - lo) / 5, n)
        out.append(np.clip(v, lo, hi).astype(dtype))
    return out


def plummer(n, dtype=np.float64, seed=42):
    """returns x,y,z and the tight bounding limits"""
    rng = np.random.default_rng(seed)
    xs = []
    have = 0
    while have < n:
        m = int((n - have) * 1.1) + 16
        R = 1.0 / np.sqrt(rng.random(m) ** (-2.0 / 3.0) - 1.0)
        u1, u2 = rng.random(m), rng.random(m)
        keep = R < 100.0
        R, u1, u2 = R[keep], u1[keep], u2[keep]
        Z = (1.0 - 2.0 * u1) * R
        th = 2 * np.pi * u2
        rxy = np.sqrt(R * R - Z * Z)
        xs.append(np.stack([rxy * np.cos(th), rxy * np.sin(th), Z]) * (3.0 * np.pi / 16.0))
        have += R.size
    p = np.concatenate(xs, axis=1)[:, :n]
    p -= p.mean(axis=1, keepdims=True)
    p = p.astype(dtype)
    lim = []
    for d in range(3):
        lim += [p[d].min(), p[d].max()]
    return [np.ascontiguousarray(p[0]), np.ascontiguousarray(p[1]), np.ascontiguousarray(p[2])], lim


def gaussian_keys(n, kdtype=np.uint64, seed=42):
    """sorted Gaussian-distributed keys in [0, 2^(3*maxLevel)) [ref: coord_samples/random.hpp:64-88]"""
    bits = 30 if np.dtype(kdtype) == np.uint32 else 63
    maxc = float(2 ** bits - 1)
    rng = np.random.default_rng(seed)
    v = rng.normal(maxc / 2, maxc / 5, n)
    bad = (v < 0) | (v > maxc)
    while bad.any():
        v[bad] = rng.normal(maxc / 2, maxc / 5, int(bad.sum()))
        bad = (v < 0) | (v > maxc)
    k = np.minimum(v, maxc * (1 - 1e-16)).astype(kdtype)
    k = np.minimum(k, np.array(2 ** bits - 1, dtype=kdtype))
    return np.sort(k)
