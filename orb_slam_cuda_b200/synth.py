"""Deterministic synthetic inputs for the ORB front end.

No datasets exist on the GPU box, so the tests and bench.py use seed-indexed textured grayscale frames that
have the shapes of the reference's example sequences (KITTI 1241x376, TUM 640x480, EuRoC 752x480; the YAMLs at
Examples/Stereo/KITTI00-02.yaml:18-19, Examples/Monocular/TUM1.yaml, Examples/Stereo/EuRoC.yaml:18-19).
Frames contain random rectangles / discs / line segments over low-frequency value noise so FAST finds corners
at every pyramid scale, and a share of flat regions so that the iniThFAST -> minThFAST fallback and empty
cells are exercised.  numpy only; nothing here touches the GPU, the oracle or the reference.
"""
from __future__ import annotations

import numpy as np

SHAPES = {"kitti": (376, 1241), "tum": (480, 640), "euroc": (480, 752)}


def _smooth(img: np.ndarray, passes: int = 1) -> np.ndarray:
    """Separable [1 2 1]/4 smoothing with edge replication."""
    out = img.astype(np.float32)
    for _ in range(passes):
        p = np.pad(out, 1, mode="edge")
        out = (p[1:-1, :-2] + 2 * p[1:-1, 1:-1] + p[1:-1, 2:]) * 0.25
        p = np.pad(out, 1, mode="edge")
        out = (p[:-2, 1:-1] + 2 * p[1:-1, 1:-1] + p[2:, 1:-1]) * 0.25
    return out


def _value_noise(rng: np.random.Generator, h: int, w: int, cell: int) -> np.ndarray:
    gh, gw = h // cell + 2, w // cell + 2
    g = rng.uniform(0.0, 1.0, size=(gh, gw)).astype(np.float32)
    ys = (np.arange(h, dtype=np.float32) / cell)
    xs = (np.arange(w, dtype=np.float32) / cell)
    y0, x0 = ys.astype(np.int32), xs.astype(np.int32)
    fy, fx = (ys - y0)[:, None], (xs - x0)[None, :]
    a = g[y0][:, x0]
    b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]
    d = g[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def textured_frame(seed: int, h: int, w: int, flat_fraction: float = 0.15) -> np.ndarray:
    """One uint8 frame; frame k of a sequence uses seed 20260000 + k."""
    rng = np.random.default_rng(seed)
    img = 40.0 + 150.0 * (0.6 * _value_noise(rng, h, w, 97) + 0.4 * _value_noise(rng, h, w, 23))
    n_shapes = int(0.0016 * h * w)
    for _ in range(n_shapes):
        kind = rng.integers(0, 3)
        cx, cy = int(rng.integers(0, w)), int(rng.integers(0, h))
        s = int(rng.integers(3, 40))
        val = float(rng.integers(0, 256))
        if kind == 0:  # rectangle
            sy = int(rng.integers(3, 40))
            img[max(cy - sy, 0):cy + sy, max(cx - s, 0):cx + s] = val
        elif kind == 1:  # disc
            y0, y1, x0, x1 = max(cy - s, 0), min(cy + s + 1, h), max(cx - s, 0), min(cx + s + 1, w)
            yy, xx = np.ogrid[y0:y1, x0:x1]
            m = (yy - cy) ** 2 + (xx - cx) ** 2 <= s * s
            img[y0:y1, x0:x1][m] = val
        else:  # triangle (half of a rectangle along the diagonal)
            y0, y1, x0, x1 = max(cy - s, 0), min(cy + s + 1, h), max(cx - s, 0), min(cx + s + 1, w)
            yy, xx = np.ogrid[y0:y1, x0:x1]
            m = (yy - cy) * (1 if rng.integers(0, 2) else -1) <= (xx - cx)
            img[y0:y1, x0:x1][m] = val
    # flat (low-texture) patches: force the minThFAST fallback / empty cells
    n_flat = max(1, int(flat_fraction * h * w / (90 * 90)))
    for _ in range(n_flat):
        cx, cy = int(rng.integers(0, w)), int(rng.integers(0, h))
        sx, sy = int(rng.integers(30, 90)), int(rng.integers(30, 90))
        base = float(rng.integers(30, 220))
        sl = (slice(max(cy - sy, 0), cy + sy), slice(max(cx - sx, 0), cx + sx))
        patch = img[sl]
        # half of the flat patches keep a faint texture (contrast ~ 10 gray levels), the others are constant
        img[sl] = base + (0.06 * (patch - patch.mean()) if rng.integers(0, 2) else 0.0)
    img = _smooth(img, 1)
    img += rng.normal(0.0, 1.5, size=(h, w)).astype(np.float32)
    return np.ascontiguousarray(np.clip(np.rint(img), 0, 255).astype(np.uint8))


def frame(kind: str, k: int) -> np.ndarray:
    h, w = SHAPES[kind]
    return textured_frame(20260000 + k, h, w)


def frame_batch(kind: str, n: int, pool: int = 32, first: int = 0) -> np.ndarray:
    """n frames [n, h, w].  `pool` distinct textures are generated; frame k beyond the pool is pool frame
    (k % pool) cyclically shifted by a k-dependent offset, so every frame is a different image while generation
    stays fast enough for 4096-frame batches."""
    h, w = SHAPES[kind]
    base = [frame(kind, first + i) for i in range(min(pool, n))]
    out = np.empty((n, h, w), dtype=np.uint8)
    for k in range(n):
        b = base[k % len(base)]
        r = k // len(base)
        out[k] = b if r == 0 else np.roll(b, (5 * r % h, 11 * r % w), axis=(0, 1))
    return out


def stereo_pair(kind: str, k: int, max_disp: int = 48):
    """Left frame and a right frame = left resampled with a smooth positive disparity field + own noise."""
    h, w = SHAPES[kind]
    left = frame(kind, k)
    rng = np.random.default_rng(77000000 + k)
    disp = 2.0 + (max_disp - 2.0) * _value_noise(rng, h, w, 160)
    xs = np.arange(w, dtype=np.float32)[None, :] + disp  # right(x) = left(x + d)
    x0 = np.clip(np.floor(xs).astype(np.int32), 0, w - 1)
    x1 = np.clip(x0 + 1, 0, w - 1)
    f = xs - np.floor(xs)
    rows = np.arange(h)[:, None]
    lf = left.astype(np.float32)
    right = lf[rows, x0] * (1 - f) + lf[rows, x1] * f
    right += rng.normal(0.0, 1.0, size=(h, w)).astype(np.float32)
    return left, np.ascontiguousarray(np.clip(np.rint(right), 0, 255).astype(np.uint8))


def shifted_frame(kind: str, k: int, dx: int = 7, dy: int = -4) -> np.ndarray:
    """Frame t+1 for frame-to-frame matching: frame k translated by (dx, dy) with fresh sensor noise."""
    a = frame(kind, k)
    rng = np.random.default_rng(88000000 + k)
    b = np.roll(a, (dy, dx), axis=(0, 1)).astype(np.float32) + rng.normal(0.0, 1.5, size=a.shape)
    return np.ascontiguousarray(np.clip(np.rint(b), 0, 255).astype(np.uint8))


def descriptor_sets(seed: int, nq: int, nd: int, dup_fraction: float = 0.5, max_flips: int = 30):
    """Random 256-bit descriptors; a share of the queries are planted near-duplicates (<= max_flips flipped
    bits) of database rows so that the TH_LOW / ratio tests fire on both sides."""
    rng = np.random.default_rng(seed)
    d = rng.integers(0, 256, size=(nd, 32), dtype=np.uint8)
    q = rng.integers(0, 256, size=(nq, 32), dtype=np.uint8)
    ndup = int(dup_fraction * nq)
    src = rng.integers(0, nd, size=ndup)
    qbits = np.unpackbits(d[src], axis=1)
    for i in range(ndup):
        nflip = int(rng.integers(0, max_flips + 1))
        pos = rng.choice(256, size=nflip, replace=False)
        qbits[i, pos] ^= 1
    q[:ndup] = np.packbits(qbits, axis=1)
    perm = rng.permutation(nq)
    return q[perm].copy(), d
