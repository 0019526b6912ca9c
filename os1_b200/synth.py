"""Seeded synthetic grayscale frames for parity tests and benchmarks.

Numpy only, so the same frames can be produced on the GPU box without cv2/scipy.
A frame is: Gaussian-smoothed (sigma 1.5) uniform noise stretched to 0..255, plus 60 constant
rectangles (strong corners), with the right-hand third squeezed to low contrast (x0.15 + 100) so that
FAST's iniTh finds nothing in some cells and the minTh retry of ORBextractor.cc:852-856 is exercised.
"""
import numpy as np


def _gauss_kernel(sigma: float) -> np.ndarray:
    r = int(np.ceil(3 * sigma))
    x = np.arange(-r, r + 1, dtype=np.float64)
    k = np.exp(-0.5 * (x / sigma) ** 2)
    return k / k.sum()


def _smooth(a: np.ndarray, sigma: float) -> np.ndarray:
    k = _gauss_kernel(sigma)
    r = len(k) // 2
    p = np.pad(a, ((0, 0), (r, r)), mode="reflect")
    out = np.zeros_like(a)
    for i, w in enumerate(k):
        out += w * p[:, i:i + a.shape[1]]
    p = np.pad(out, ((r, r), (0, 0)), mode="reflect")
    out2 = np.zeros_like(a)
    for i, w in enumerate(k):
        out2 += w * p[i:i + a.shape[0], :]
    return out2


def scene(height: int, width: int, seed: int = 0, margin: int = 0) -> np.ndarray:
    """One static scene, (height+2*margin) x (width+2*margin) uint8."""
    H, W = height + 2 * margin, width + 2 * margin
    rng = np.random.default_rng(seed)
    a = _smooth(rng.random((H, W)), 1.5)
    a = (a - a.min()) / (a.max() - a.min()) * 255.0
    for _ in range(60):
        x0 = int(rng.integers(0, W - 8)); y0 = int(rng.integers(0, H - 8))
        w = int(rng.integers(8, max(9, W // 6))); h = int(rng.integers(8, max(9, H // 6)))
        a[y0:y0 + h, x0:x0 + w] = float(rng.integers(0, 256))
    t = (2 * W) // 3
    a[:, t:] = a[:, t:] * 0.15 + 100.0
    return np.clip(np.rint(a), 0, 255).astype(np.uint8)


def frame(height: int, width: int, seed: int = 0) -> np.ndarray:
    return scene(height, width, seed)


def video(nframes: int, height: int, width: int, seed: int = 0, max_shift: int = 24) -> np.ndarray:
    """nframes x height x width uint8: one scene seen through a window moved by a seeded integer shift per frame."""
    sc = scene(height, width, seed, margin=max_shift)
    rng = np.random.default_rng(seed + 7919)
    out = np.empty((nframes, height, width), np.uint8)
    for f in range(nframes):
        dx = int(rng.integers(0, 2 * max_shift + 1)); dy = int(rng.integers(0, 2 * max_shift + 1))
        out[f] = sc[dy:dy + height, dx:dx + width]
    return out
