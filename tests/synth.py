"""Seeded synthetic stereo frames: smooth background + rectangles/discs + pixel noise.

Used by tests (small counts), bench.py (batches) -- pure numpy so it runs on the GPU box without /root/reference.
"""
import numpy as np


def _value_noise(rng, h, w, octaves=3, amp=60.0):
    out = np.zeros((h, w), np.float32)
    for o in range(octaves):
        n = 4 << o
        g = rng.random((n + 2, n + 2), dtype=np.float32)
        ys = np.linspace(0, n, h, endpoint=False, dtype=np.float32)
        xs = np.linspace(0, n, w, endpoint=False, dtype=np.float32)
        y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
        fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
        a = g[y0][:, x0]; b = g[y0][:, x0 + 1]; c = g[y0 + 1][:, x0]; d = g[y0 + 1][:, x0 + 1]
        out += (amp / (1 << o)) * ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy)
    return out


def synth_left(seed, w=752, h=480, nshapes=400):
    rng = np.random.default_rng(1234 + seed)
    img = 40.0 + _value_noise(rng, h, w)
    for _ in range(nshapes):
        cx, cy = int(rng.integers(0, w)), int(rng.integers(0, h))
        s = int(rng.integers(4, 40))
        g = float(rng.integers(0, 256))
        kind, t, frac = rng.random(), rng.random() * np.pi, rng.random()
        x0, x1, y0, y1 = max(cx - s, 0), min(cx + s + 1, w), max(cy - s, 0), min(cy + s + 1, h)
        yy, xx = np.mgrid[y0:y1, x0:x1]
        if kind < 0.3:
            m = (xx - cx) ** 2 + (yy - cy) ** 2 <= (s // 2) ** 2
        elif kind < 0.65:
            u = (xx - cx) * np.cos(t) + (yy - cy) * np.sin(t)
            v = -(xx - cx) * np.sin(t) + (yy - cy) * np.cos(t)
            m = (np.abs(u) <= s // 2) & (np.abs(v) <= max(2, s // 3))
        else:
            m = (np.abs(xx - cx) <= s // 2) & (np.abs(yy - cy) <= max(2, int(s * frac)))
        img[y0:y1, x0:x1][m] = g
    return img, rng


def synth_stereo(seed, w=752, h=480, nshapes=400, dmin=2.0, dmax=60.0):
    """Rectified pair: right = left warped horizontally by a smooth disparity field, + independent noise."""
    base, rng = synth_left(seed, w, h, nshapes)
    disp = dmin + (dmax - dmin) * (_value_noise(rng, h, w, octaves=2, amp=1.0) / 1.5)
    xs = np.arange(w, dtype=np.float32)[None, :] + disp       # right(x) = left(x + d)
    x0 = np.clip(np.floor(xs).astype(np.int32), 0, w - 1)
    x1 = np.clip(x0 + 1, 0, w - 1)
    f = xs - np.floor(xs)
    rows = np.arange(h)[:, None]
    right = base[rows, x0] * (1 - f) + base[rows, x1] * f
    left = base + rng.normal(0, 2.0, base.shape)
    right = right + rng.normal(0, 2.0, base.shape)
    return (np.clip(np.rint(left), 0, 255).astype(np.uint8), np.clip(np.rint(right), 0, 255).astype(np.uint8))


def rectify_maps(w, h, k1=-0.28, k2=0.07, tilt=0.012, shift=(3.25, -2.5), scale=1.04):
    """Deterministic float32 rectification maps (M1 = x, M2 = y) of a radial-distortion + small-rotation camera, using
    only correctly-rounded float32 operations (+ - * /), so every machine computes identical bits.  Parts of the
    destination map outside the source on purpose (constant-0 border of cv::remap)."""
    f = np.float32
    yy, xx = np.meshgrid(np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32), indexing="ij")
    cx, cy, fl = f(w / 2 - 0.5), f(h / 2 - 0.5), f(0.6 * w)
    x = (xx - cx) / fl * f(scale); y = (yy - cy) / fl * f(scale)
    xr = x - f(tilt) * y; yr = y + f(tilt) * x
    r2 = xr * xr + yr * yr
    d = f(1) + f(k1) * r2 + f(k2) * r2 * r2
    return (xr * d * fl + cx + f(shift[0])).astype(np.float32), (yr * d * fl + cy + f(shift[1])).astype(np.float32)
