"""Shared helpers for the parity tests."""
import hashlib

import numpy as np


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def pairs_hash(pairs):
    h = 1469598103934665603
    for v in np.asarray(pairs).reshape(-1).tolist():
        h = ((h ^ int(v)) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h


def tform_vec(d):
    """dict from transform_to_dict / lrb -> flat vector in the golden layout."""
    return np.array([d["width"], d["height"], *d["top_left"], *d["top_right"], *d["bottom_left"], *d["bottom_right"],
                     *d["horizontal_vp"], *d["vertical_vp"]], dtype=np.float64)


def seeds_equal_modulo_ties(seeds_a, seeds_b, mag):
    """Same seed set, same order except inside groups of exactly equal magnitude (std::sort is unstable there)."""
    if len(seeds_a) != len(seeds_b):
        return False
    va = mag[seeds_a[:, 0], seeds_a[:, 1]]
    vb = mag[seeds_b[:, 0], seeds_b[:, 1]]
    if not np.array_equal(va, vb):
        return False
    if np.any(np.diff(va) > 0):
        return False
    ka = np.lexsort((seeds_a[:, 1], seeds_a[:, 0], -va.astype(np.float64)))
    kb = np.lexsort((seeds_b[:, 1], seeds_b[:, 0], -vb.astype(np.float64)))
    return np.array_equal(seeds_a[ka], seeds_b[kb])


def n_value_ties(seeds, mag):
    v = mag[seeds[:, 0], seeds[:, 1]]
    return int(np.sum(np.diff(v) == 0))


def labels_equal_up_to_relabelling(la, lb):
    """Two label maps (-1 = background) describe the same partition."""
    if not np.array_equal(la >= 0, lb >= 0):
        return False
    m = la >= 0
    a = la[m].astype(np.int64); b = lb[m].astype(np.int64)
    if a.size == 0:
        return True
    pa = np.unique(np.stack([a, b], 1), axis=0)
    return len(pa) == len(np.unique(a)) == len(np.unique(b))


# sin / cos of float(i*pi)/8 as float32 (pinned by the bit-exact grad8 digests in tests/golden)
SIN8 = np.array([float.fromhex(x) for x in ("0x0p+0", "0x1.87de2cp-2", "0x1.6a09e6p-1", "0x1.d906bcp-1", "0x1p+0",
                                            "0x1.d906bep-1", "0x1.6a09e6p-1", "0x1.87de2ep-2")], dtype=np.float32)
COS8 = np.array([float.fromhex(x) for x in ("0x1p+0", "0x1.d906bcp-1", "0x1.6a09e6p-1", "0x1.87de2ap-2", "-0x1.777a5cp-25",
                                            "-0x1.87de28p-2", "-0x1.6a09e6p-1", "-0x1.d906bcp-1")], dtype=np.float32)
