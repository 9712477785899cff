"""Seeded synthetic inputs for the SIFT hot path.

Frames are 8-bit-valued float32 images (0..255, integers) so the reference's thresholds
(`Detector.cu:362`, `SiftOps.cu:90`) behave as on real camera frames; descriptor sets mimic the
reference's descriptor encoding (`SiftOps.cu:606-621`: L2-normalise, clip 0.2, renormalise, x512,
store as half). Pure numpy, no library-dependent resampling, so the same seed gives the same
bytes everywhere (dev container, GPU box, judge).
"""
from __future__ import annotations

import numpy as np

FRAME_SEED = 1234
QUERY_SEED = 42
TRAIN_SEED = 43


def _cubic_upsample(grid: np.ndarray, cell: int, h: int, w: int) -> np.ndarray:
    """Separable Catmull-Rom upsampling of a coarse noise grid by `cell`."""

    def axis_weights(n_out: int):
        u = (np.arange(n_out, dtype=np.float64) + 0.5) / cell + 0.5
        i = np.floor(u).astype(np.int64)
        t = u - i
        w0 = ((-t + 2.0) * t - 1.0) * t * 0.5
        w1 = ((3.0 * t - 5.0) * t * t + 2.0) * 0.5
        w2 = ((-3.0 * t + 4.0) * t + 1.0) * t * 0.5
        w3 = ((t - 1.0) * t * t) * 0.5
        return i, np.stack([w0, w1, w2, w3], axis=0)

    iy, wy = axis_weights(h)
    ix, wx = axis_weights(w)
    rows = sum(wy[k][:, None] * grid[iy + k, :] for k in range(4))
    return sum(wx[k][None, :] * rows[:, ix + k] for k in range(4))


def synth_frame(rows: int, cols: int, frame_index: int = 0, density: float = 1.0) -> np.ndarray:
    """Procedurally textured frame: 6 octaves of smooth noise + 64 random discs / rectangles.

    `density` scales the amplitude of the finest noise octaves (more fine texture -> more
    keypoints); 1.0 is the default workload of bench.py and the parity tests.
    """
    rng = np.random.Generator(np.random.PCG64(FRAME_SEED + frame_index))
    img = np.zeros((rows, cols), dtype=np.float64)
    for k in range(6):
        cell = 2 ** k
        gh, gw = -(-rows // cell) + 5, -(-cols // cell) + 5
        grid = rng.random((gh, gw))
        amp = float(cell) if k >= 2 else float(cell) * density
        img += amp * _cubic_upsample(grid, cell, rows, cols) if cell > 1 else amp * grid[2:2 + rows, 2:2 + cols]
    yy, xx = np.mgrid[0:rows, 0:cols]
    span = img.max() - img.min()
    for _ in range(64):
        level = img.min() + span * rng.random()
        cy, cx = rng.random() * rows, rng.random() * cols
        if rng.random() < 0.5:
            rad = 4 + rng.random() * 0.06 * min(rows, cols)
            mask = (yy - cy) ** 2 + (xx - cx) ** 2 <= rad * rad
        else:
            hh, hw = 3 + rng.random() * 0.06 * rows, 3 + rng.random() * 0.06 * cols
            mask = (np.abs(yy - cy) <= hh) & (np.abs(xx - cx) <= hw)
        img[mask] = 0.5 * img[mask] + 0.5 * level
    img = (img - img.min()) / (img.max() - img.min()) * 255.0
    return np.floor(img).astype(np.float32)


def next_frame(prev: np.ndarray, frame_index: int) -> np.ndarray:
    """'Consecutive frame': previous frame translated 2 px, rotated 0.5 deg (bilinear), + noise."""
    rows, cols = prev.shape
    rng = np.random.Generator(np.random.PCG64(FRAME_SEED + 100000 + frame_index))
    th = np.deg2rad(0.5)
    yy, xx = np.mgrid[0:rows, 0:cols].astype(np.float64)
    cy, cx = rows / 2.0, cols / 2.0
    sx = np.cos(th) * (xx - cx) + np.sin(th) * (yy - cy) + cx - 2.0
    sy = -np.sin(th) * (xx - cx) + np.cos(th) * (yy - cy) + cy - 2.0
    sx = np.clip(sx, 0, cols - 1.001)
    sy = np.clip(sy, 0, rows - 1.001)
    x0, y0 = np.floor(sx).astype(np.int64), np.floor(sy).astype(np.int64)
    fx, fy = sx - x0, sy - y0
    p = prev.astype(np.float64)
    out = (p[y0, x0] * (1 - fx) * (1 - fy) + p[y0, x0 + 1] * fx * (1 - fy)
           + p[y0 + 1, x0] * (1 - fx) * fy + p[y0 + 1, x0 + 1] * fx * fy)
    out += rng.normal(0.0, 2.0, size=out.shape)
    return np.floor(np.clip(out, 0, 255)).astype(np.float32)


def blocks_frame(rows: int, cols: int, seed: int) -> np.ndarray:
    """Axis-aligned rectangles of random grey levels: gradients point exactly along the axes, i.e. the orientation
    samples sit ON histogram bin edges (0 / 90 / 180 / 270 degrees) and the descriptor samples where the vote is
    discontinuous in the gradient angle -- the cases the fast paths of orientation_kernel / descriptor_kernel must
    hand to the reference's own expressions."""
    rng = np.random.default_rng(seed)
    img = np.full((rows, cols), 96, np.float32)
    for _ in range(260):
        h, w = rng.integers(6, 60, 2)
        y, x = rng.integers(0, rows - h), rng.integers(0, cols - w)
        img[y:y + h, x:x + w] = float(rng.integers(0, 256))
    return img


def _encode(v: np.ndarray) -> np.ndarray:
    v = v / np.maximum(np.linalg.norm(v, axis=1, keepdims=True), 1e-7)
    v = np.minimum(v, 0.2)
    v = v / np.maximum(np.linalg.norm(v, axis=1, keepdims=True), 1e-7)
    return (np.clip(v, 0, 1) * 512.0).astype(np.float16)


def synth_descriptors(n_query: int, n_train: int, dim: int = 128):
    """(query, train) float16 [n,128] descriptor sets; ~60% of train rows are noisy copies of
    (permuted) query rows so the ratio test accepts a realistic share of matches."""
    rq = np.random.Generator(np.random.PCG64(QUERY_SEED))
    rt = np.random.Generator(np.random.PCG64(TRAIN_SEED))
    q = np.abs(rq.standard_normal((n_query, dim)))
    q[rq.random((n_query, dim)) < 0.3] = 0.0
    query = _encode(q)
    src = rt.integers(0, n_query, size=n_train)
    base = query[src].astype(np.float64)
    copy = rt.random(n_train) < 0.6
    noisy = np.abs(base + rt.normal(0.0, 8.0, size=base.shape))
    fresh = np.abs(rt.standard_normal((n_train, dim)))
    fresh[rt.random((n_train, dim)) < 0.3] = 0.0
    train = _encode(np.where(copy[:, None], noisy, fresh))
    return query, train
