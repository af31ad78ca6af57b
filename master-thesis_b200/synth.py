"""Seeded synthetic rectified stereo pairs; numpy only, so inputs are identical wherever they are
generated.  Input synthesis for tests, the golden generator and bench.py -- not part of the compute path."""
import numpy as np


def _blur121(a):
    """separable 5-tap binomial blur (sigma ~ 1.2) in integer arithmetic -> deterministic across numpy builds."""
    k = np.array([1, 4, 6, 4, 1], np.int32)
    a = a.astype(np.int32)
    p = np.pad(a, ((0, 0), (2, 2)), mode="edge")
    h = sum(k[i] * p[:, i:i + a.shape[1]] for i in range(5))
    p = np.pad(h, ((2, 2), (0, 0)), mode="edge")
    v = sum(k[i] * p[i:i + a.shape[0], :] for i in range(5))
    return ((v + 128) >> 8).astype(np.uint8)


def synth_pair(H, W, D, seed, color=False):
    """Returns (left, right) uint8 gray (H, W), or BGR (H, W, 3) when color=True.
    Ground-truth disparity d(y,x) = floor(0.15 D + 0.5 D y/H) + D/8 inside the central third;
    right(y, x) = left(y, min(x + d, W-1))."""
    rng = np.random.default_rng(seed)
    T = _blur121(rng.integers(0, 256, size=(H, W + D), dtype=np.uint8))
    left = T[:, D:D + W]
    yy, xx = np.mgrid[0:H, 0:W]
    d = np.floor(0.15 * D + 0.5 * D * yy / H).astype(np.int64)
    box = (yy >= H // 3) & (yy < 2 * H // 3) & (xx >= W // 3) & (xx < 2 * W // 3)
    d = d + np.where(box, D // 8, 0)
    xs = np.minimum(xx + d, W - 1)
    right = np.take_along_axis(left, xs, axis=1)
    left = np.ascontiguousarray(left)
    right = np.ascontiguousarray(right)
    if not color:
        return left, right
    n = rng.integers(-3, 4, size=(2, H, W, 3))
    lb = np.clip(left[..., None].astype(np.int64) + n[0], 0, 255).astype(np.uint8)
    rb = np.clip(right[..., None].astype(np.int64) + n[1], 0, 255).astype(np.uint8)
    return lb, rb


def satdrift_pair(H=80, W=260, D=32, seed=5):
    """Adversarial pair for the int16 saturation regime (block_size 27): the top half is flat 255 (left) against
    flat 0 (right) so the raw-intensity BT cost is 63 everywhere and the vertical block sum sticks at 32767; the
    bottom half is an ordinary textured pair, so the recurrence leaves saturation and drifts."""
    l, r = synth_pair(H, W, D, seed)
    l = l.copy()
    r = r.copy()
    l[:H // 2] = 255
    r[:H // 2] = 0
    return l, r
