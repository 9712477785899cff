"""Comparators implementing the tolerances of BASELINE.json.north_star (test infrastructure)."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def record(tag, info):
    """Appends one JSON line with the measured parity figures of a comparison (excluded-by-padding fraction,
    descriptor error quantiles, set agreement ...) to gpurun_out/parity_metrics.jsonl, the file that comes back
    from the GPU box; profiles/r2_parity_metrics.jsonl is a committed copy of it."""
    try:
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "parity_metrics.jsonl"), "a") as f:
            f.write(json.dumps({"tag": tag, **{k: (v.item() if hasattr(v, "item") else v) for k, v in info.items()}}) + "\n")
    except OSError:
        pass


def keys(kpts, feats):
    """Exact-valued identity of a returned keypoint: x, y (integer px * 2^o, Q10), layer, packed octave,
    angle (360 - 10*bin, Q9)."""
    return [tuple(r) for r in np.concatenate([kpts[:, :3], feats[:, [0, 3]]], axis=1).tolist()]


def set_agreement(k1, f1, k2, f2):
    from collections import Counter

    a, b = Counter(keys(k1, f1)), Counter(keys(k2, f2))
    inter = sum((a & b).values())
    return inter / max(len(k1), len(k2), 1)


def rows_agreement(a, b):
    """Multiset agreement of the rows of two 2-D float arrays (bit patterns), |a & b| / max(|a|, |b|)."""
    from collections import Counter

    ca = Counter(map(bytes, np.ascontiguousarray(a, np.float32)))
    cb = Counter(map(bytes, np.ascontiguousarray(b, np.float32)))
    return sum((ca & cb).values()) / max(len(a), len(b), 1)


def paired(k1, f1, d1, k2, f2, d2):
    """Pairs entries with the same key (first occurrence each) -> index arrays into 1 and 2."""
    pos = {}
    for i, k in enumerate(keys(k2, f2)):
        pos.setdefault(k, []).append(i)
    i1, i2 = [], []
    for i, k in enumerate(keys(k1, f1)):
        lst = pos.get(k)
        if lst:
            i1.append(i)
            i2.append(lst.pop(0))
    return np.array(i1, int), np.array(i2, int)


def desc_l2(d1, d2):
    """L2 distance between normalised descriptors (stored scale 512 -> unit scale)."""
    return np.linalg.norm((d1.astype(np.float32) - d2.astype(np.float32)) / 512.0, axis=1)


def _may_touch(kpts, feats, octaves):
    """Cheap superset: the bounding patch of the descriptor reaches the last column of a padded octave."""
    octv = feats[:, 0].astype(np.int64) & 255
    scale = 1.0 / (2.0 ** octv)
    x = np.rint(kpts[:, 0] * scale)
    radius = np.rint(3.0 * (feats[:, 1] * scale * 0.5) * 1.4142135623730951 * 2.5)
    w = np.array([octaves[o][0] for o in octv])
    pitch = np.array([octaves[o][2] for o in octv])
    return (pitch > w) & (x + radius >= w - 1)


def touches_padding(kpts, feats, octaves, margin=1e-3):
    """True for keypoints with at least one ACCEPTED descriptor sample in the last column (px == w-1) of an octave
    whose rows are padded (pitch > w). There the reference's gradient reads one float past the row (quirk Q12,
    SiftOps.cu:553-557), i.e. UNINITIALISED cudaMallocPitch padding: its result depends on what the allocation held
    before, so those descriptors cannot be pinned (the engine and the CPU oracle read zeros there). The test
    restates the sample geometry of SiftOps.cu:466-553 per keypoint: a sample (i, j) of the (2*radius+1)^2 patch is
    accepted when its rotated window coordinates rbin, cbin lie in (-1, 4) and the pixel strictly inside the image;
    `margin` widens the window test by 1e-3 bins so that float-rounding differences cannot hide a sample."""
    n = len(kpts)
    out = np.zeros(n, bool)
    for k in np.nonzero(_may_touch(kpts, feats, octaves))[0]:
        packed = int(feats[k, 0])
        octv = packed & 255
        w, h = int(octaves[octv][0]), int(octaves[octv][1])
        scale = np.float32(1.0 / (1 << octv))
        ix, iy = int(np.rint(np.float32(kpts[k, 0]) * scale)), int(np.rint(np.float32(kpts[k, 1]) * scale))
        angle = np.float32(360.0) - np.float32(feats[k, 3])
        if abs(angle - 360.0) < 1e-6:
            angle = np.float32(0.0)
        scl = np.float32(feats[k, 1]) * scale * np.float32(0.5)
        hist_width = np.float32(3.0) * scl
        radius = int(np.rint(hist_width * np.float32(1.4142135623730951) * np.float32(5.0) * np.float32(0.5)))
        radius = int(min(radius, np.sqrt(float(w) * w + float(h) * h)))
        j = w - 1 - ix                      # the only patch column that lands on px == w-1
        if j < -radius or j > radius:
            continue
        i = np.arange(-radius, radius + 1, dtype=np.float32)
        cos_t = np.cos(np.deg2rad(angle)) / hist_width
        sin_t = np.sin(np.deg2rad(angle)) / hist_width
        cbin = j * cos_t - i * sin_t + 1.5
        rbin = j * sin_t + i * cos_t + 1.5
        py = iy + i
        acc = (rbin > -1 - margin) & (rbin < 4 + margin) & (cbin > -1 - margin) & (cbin < 4 + margin) & (py > 0) & (py < h)
        out[k] = bool(acc.any()) and (w - 1 > 0)
    return out


def check_results(got, want, *, min_agree=0.995, desc_tol=1e-3, desc_frac=1.0, size_rtol=1e-3, octaves=None,
                  max_undefined=0.25, tag=None):
    """got / want: dict(n, kpts, feats, desc). Returns a dict of measured figures, asserts the bars:
    >= 99.5% keypoint-set agreement; on the paired keypoints x/y/layer/octave/angle are identical
    by construction of the key, size within 1e-3 relative, |contrast| within 1e-5, and normalised descriptors
    within `desc_tol` L2 for `desc_frac` (default: ALL) of the paired keypoints whose reference descriptor is
    defined (see touches_padding; the excluded fraction is recorded and capped by `max_undefined`)."""
    n1, n2 = got["n"], want["n"]
    k1, f1, d1 = got["kpts"][:n1], got["feats"][:n1], got["desc"][:n1]
    k2, f2, d2 = want["kpts"][:n2], want["feats"][:n2], want["desc"][:n2]
    agree = set_agreement(k1, f1, k2, f2)
    i1, i2 = paired(k1, f1, d1, k2, f2, d2)
    out = dict(n_got=n1, n_want=n2, agree=agree, paired=len(i1))
    assert agree >= min_agree, out
    if len(i1):
        size_err = np.abs(f1[i1, 1] - f2[i2, 1]) / np.maximum(np.abs(f2[i2, 1]), 1e-12)
        con_err = np.abs(f1[i1, 2] - f2[i2, 2])
        l2 = desc_l2(d1[i1], d2[i2])
        if octaves is not None:  # descriptors that depend on uninitialised padding in the reference
            undefined = touches_padding(k2[i2], f2[i2], octaves)
            out.update(undefined_frac=float(undefined.mean()), undefined_n=int(undefined.sum()),
                       may_touch_frac=float(_may_touch(k2[i2], f2[i2], octaves).mean()))
            if undefined.any():
                out.update(undefined_l2_max=float(l2[undefined].max()))
            l2 = l2[~undefined]
        if len(l2):
            out.update(desc_l2_max=float(l2.max()), desc_l2_p99=float(np.quantile(l2, 0.99)),
                       desc_within=float((l2 <= desc_tol).mean()), desc_compared=int(len(l2)))
        out.update(size_rel_max=float(size_err.max()), contrast_abs_max=float(con_err.max()))
        if tag:
            record(tag, out)
        if octaves is not None:
            assert out["undefined_frac"] <= max_undefined, out
        assert size_err.max() <= size_rtol, out
        assert con_err.max() <= 1e-5, out
        if len(l2):
            assert (l2 <= desc_tol).mean() >= desc_frac, out
    elif tag:
        record(tag, out)
    return out
