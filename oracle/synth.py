"""Synthetic surface-sampled voxel clouds.  Test/bench input generator.

Pure NumPy, seeded.  Shapes:
  block128   one sphere R=36 in a 128^3 cube               (~22 k voxels)
  frame10    four spheres stacked in a 1024^3 grid         (~818 k voxels, "longdress-shaped")
  frame11    the same x1.75 in a 2048^3 grid               (~2.5 M voxels, "Owlii-shaped")
Colours: smooth sinusoid field + noise, rounded to k/255 (8-bit), as floats in [0,1].
"""
import numpy as np


def _sphere_surface(rng, radius, centre, n_samples):
    v = rng.standard_normal((n_samples, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    p = v * radius + np.asarray(centre, dtype=np.float64)
    return np.floor(p).astype(np.int32)


def _dedupe(p):
    # first-occurrence unique, keeps the sampling order (not sorted)
    key = (p[:, 0].astype(np.int64) << 42) | (p[:, 1].astype(np.int64) << 21) | p[:, 2].astype(np.int64)
    _, first = np.unique(key, return_index=True)
    first.sort()
    return p[first]


def colours_for(points, seed=0):
    rng = np.random.default_rng(seed + 1000)
    A = np.array([[1.0, 0.5, 0.25], [0.3, 1.0, 0.6], [0.7, 0.2, 1.0]])
    phi = np.array([0.0, 2.0, 4.0])
    f = 0.5 + 0.5 * np.sin(2 * np.pi * (points.astype(np.float64) @ A.T) / 256.0 + phi)
    f += rng.normal(0.0, 0.02, size=f.shape)
    f = np.clip(f, 0.0, 1.0)
    return (np.round(f * 255.0) / 255.0).astype(np.float32)


def block128(seed=0, n_samples=400_000):
    rng = np.random.default_rng(seed)
    p = _dedupe(_sphere_surface(rng, 36.0, (64, 64, 64), n_samples))
    return p, colours_for(p, seed)


def _stack(seed, scale, samples, jitter=0):
    rng = np.random.default_rng(seed)
    radii = np.array([150, 110, 90, 70]) * scale
    ys = np.array([300, 560, 760, 910]) * scale
    c0 = 512 * scale
    j = rng.integers(-jitter, jitter + 1, size=3) if jitter else np.zeros(3, dtype=np.int64)
    parts = []
    for r, y, n in zip(radii, ys, samples):
        parts.append(_sphere_surface(rng, float(r), (c0 + j[0], y + j[1], c0 + j[2]), int(n)))
    p = _dedupe(np.concatenate(parts, axis=0))
    return p, colours_for(p, seed)


def frame10(seed=0, jitter=0):
    """10-bit longdress-shaped frame (~818 k voxels)."""
    return _stack(seed, 1.0, (6e6, 4e6, 3e6, 2e6), jitter)


def frame11(seed=0, jitter=8):
    """11-bit Owlii-shaped frame (~2.5 M voxels)."""
    return _stack(seed, 1.75, (20e6, 12e6, 9e6, 6e6), jitter)


def small_cloud(seed=0, radius=9.0, centre=(16, 16, 16), n_samples=6000):
    """Tiny sphere for unit tests (fits a 32^3 dense grid)."""
    rng = np.random.default_rng(seed)
    p = _dedupe(_sphere_surface(rng, radius, centre, n_samples))
    return p, colours_for(p, seed)


def random_cloud(seed, n, extent, batch=1):
    """Uniform random unique voxels in [0,extent)^3, with a batch column -> (N,4) int32."""
    rng = np.random.default_rng(seed)
    out = []
    for b in range(batch):
        p = _dedupe(rng.integers(0, extent, size=(n, 3)).astype(np.int32))
        out.append(np.concatenate([np.full((len(p), 1), b, np.int32), p], axis=1))
    return np.concatenate(out, axis=0)
