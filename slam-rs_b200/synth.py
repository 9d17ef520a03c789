"""Deterministic synthetic inputs for the benchmark and the tests.  Standalone: numpy only.

The generators avoid cv2 so that fixtures are reproducible anywhere numpy runs: the Gaussian
smoothing of the noise texture is a separable float64 numpy convolution.
"""
import numpy as np


def _gauss_kernel(sigma):
    r = int(np.ceil(3 * sigma))
    x = np.arange(-r, r + 1, dtype=np.float64)
    k = np.exp(-(x * x) / (2 * sigma * sigma))
    return k / k.sum()


def _blur(a, sigma):
    k = _gauss_kernel(sigma)
    r = len(k) // 2
    p = np.pad(a, ((0, 0), (r, r)), mode="reflect")
    out = np.zeros_like(a)
    for i, kv in enumerate(k):
        out += kv * p[:, i:i + a.shape[1]]
    p = np.pad(out, ((r, r), (0, 0)), mode="reflect")
    out2 = np.zeros_like(a)
    for i, kv in enumerate(k):
        out2 += kv * p[i:i + a.shape[0], :]
    return out2


def frame(h, w, seed, sparse=False):
    """Dense-texture grayscale frame: blurred uniform noise, min-max normalised, plus 200 random
    squares blended at 50 %.  ``sparse=True`` gives the 8x-upsampled-noise + 400 flat squares
    sensitivity variant (fewer FAST corners)."""
    rng = np.random.default_rng(seed)
    if not sparse:
        n = rng.integers(0, 256, (h, w)).astype(np.float64)
        n = _blur(n, 1.5)
        n = (n - n.min()) / (n.max() - n.min()) * 255.0
        nsq, blend = 200, True
    else:
        hs, ws = (h + 7) // 8, (w + 7) // 8
        n = rng.integers(0, 256, (hs, ws)).astype(np.float64)
        n = np.kron(n, np.ones((8, 8)))[:h, :w]
        n = _blur(n, 1.0)
        nsq, blend = 400, False
    for _ in range(nsq):
        s = int(rng.integers(8, 40))
        y = int(rng.integers(0, max(h - s, 1)))
        x = int(rng.integers(0, max(w - s, 1)))
        v = float(rng.integers(0, 256))
        if blend:
            n[y:y + s, x:x + s] = 0.5 * n[y:y + s, x:x + s] + 0.5 * v
        else:
            n[y:y + s, x:x + s] = v
    return np.clip(np.rint(n), 0, 255).astype(np.uint8)


def sequence(h, w, nframes, canvas_seed=100, shift=3, noise_sigma=2.0, noise_seed0=1000):
    """KITTI-shaped sequence (config C2): frame t is a window of a wide canvas shifted ``shift`` px
    per frame plus N(0, sigma^2) sensor noise.  Returns uint8 [nframes, h, w]."""
    canvas = frame(h, w + shift * nframes, canvas_seed)
    out = np.empty((nframes, h, w), np.uint8)
    for t in range(nframes):
        win = canvas[:, shift * t: shift * t + w].astype(np.float32)
        rng = np.random.default_rng(noise_seed0 + t)
        win += np.float32(noise_sigma) * rng.standard_normal(win.shape, dtype=np.float32)
        out[t] = np.clip(np.rint(win), 0, 255).astype(np.uint8)
    return out


def random_descriptors(n, seed):
    return np.random.default_rng(seed).integers(0, 256, (n, 32), dtype=np.uint8)


def map_and_frame_descriptors(n_map, n_frame, f=0, map_seed=7):
    """Config C4: ``n_frame/2`` rows are noisy copies (Binomial(256, 0.1) bit flips) of map rows,
    the rest are fresh random rows."""
    m = random_descriptors(n_map, map_seed)
    rng = np.random.default_rng(8 + f)
    half = n_frame // 2
    pick = rng.choice(n_map, half, replace=False)
    flips = rng.random((half, 256)) < 0.1
    noisy = m[pick] ^ np.packbits(flips, axis=1, bitorder="little")
    fresh = rng.integers(0, 256, (n_frame - half, 32), dtype=np.uint8)
    return m, np.concatenate([noisy, fresh]).astype(np.uint8)
