"""Comparators implementing the tolerances of BASELINE.json.north_star (test infrastructure)."""
import numpy as np


def keys(kpts, feats):
    """Exact-valued identity of a returned keypoint: x, y (integer px * 2^o, Q10), layer, packed octave,
    angle (360 - 10*bin, Q9)."""
    return [tuple(r) for r in np.concatenate([kpts[:, :3], feats[:, [0, 3]]], axis=1).tolist()]


def set_agreement(k1, f1, k2, f2):
    from collections import Counter

    a, b = Counter(keys(k1, f1)), Counter(keys(k2, f2))
    inter = sum((a & b).values())
    return inter / max(len(k1), len(k2), 1)


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


def touches_padding(kpts, feats, octaves):
    """True for keypoints whose descriptor patch reaches the last column of an octave whose rows are padded
    (pitch > width). There the reference reads one float past the row (quirk Q12, SiftOps.cu:553-557), i.e.
    UNINITIALISED cudaMallocPitch padding: its result is undefined (it depends on what the allocation held
    before), so those descriptors cannot be pinned. The engine and the CPU oracle read zeros there."""
    octv = feats[:, 0].astype(np.int64) & 255
    scale = 1.0 / (2.0 ** octv)
    x = np.rint(kpts[:, 0] * scale)
    radius = np.rint(3.0 * (feats[:, 1] * scale * 0.5) * 1.4142135623730951 * 2.5)
    w = np.array([octaves[o][0] for o in octv])
    pitch = np.array([octaves[o][2] for o in octv])
    return (pitch > w) & (x + radius >= w - 1)


def check_results(got, want, *, min_agree=0.995, desc_tol=1e-3, desc_frac=0.99, size_rtol=1e-3, octaves=None):
    """got / want: dict(n, kpts, feats, desc). Returns a dict of measured figures, asserts the bars:
    >= 99.5% keypoint-set agreement; on the paired keypoints x/y/layer/octave/angle are identical
    by construction of the key, size within 1e-3 relative, |contrast| within 1e-5, and descriptors
    within `desc_tol` L2 for at least `desc_frac` of the keypoints (the reference itself is not
    run-to-run deterministic there: float atomics, tests/golden/envelope_752x480.npz)."""
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
            out.update(undefined_frac=float(undefined.mean()))
            assert undefined.mean() <= 0.5, out
            l2 = l2[~undefined]
        out.update(size_rel_max=float(size_err.max()), contrast_abs_max=float(con_err.max()),
                   desc_l2_max=float(l2.max()), desc_l2_p99=float(np.quantile(l2, 0.99)),
                   desc_within=float((l2 <= desc_tol).mean()))
        assert size_err.max() <= size_rtol, out
        assert con_err.max() <= 1e-5, out
        assert (l2 <= desc_tol).mean() >= desc_frac, out
    return out
