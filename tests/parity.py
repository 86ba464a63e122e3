"""Shared parity helpers for the tests (checker side only)."""
import glob
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(prefix):
    files = sorted(glob.glob(os.path.join(GOLDEN, prefix + "_*.npz")))
    assert files, f"no golden files for {prefix}"
    return [(os.path.basename(f)[:-4], dict(np.load(f))) for f in files]


def assert_close(a, b, tol=1e-4, what=""):
    """|a-b| <= tol * max(1e-30, max|b|) elementwise-relative to the tensor scale plus rtol on the element
    (north_star: 1e-4 relative error in fp32)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    scale = max(float(np.max(np.abs(b))) if b.size else 0.0, 1e-30)
    err = np.abs(a - b)
    bound = tol * np.maximum(np.abs(b), 0.05 * scale)
    bad = err > bound
    assert not bad.any(), f"{what}: {bad.sum()} / {bad.size} outside tol {tol}; max err {err.max():.3e} (scale {scale:.3e})"


def knn_sets_equivalent(idx_a, idx_b, xyz, query, what=""):
    """idx_a (ours/canonical) vs idx_b (reference _knn_torch): equal as sets per query, after
    (1) collapsing exact ties and (2) adjudicating residual mismatches with float64 distances in
    favour of idx_a (the reference's cdist uses the |x|^2+|y|^2-2xy expansion)."""
    B, M, k = idx_a.shape
    n_mismatch = 0
    for b in range(B):
        for m in range(M):
            sa, sb = set(idx_a[b, m].tolist()), set(idx_b[b, m].tolist())
            if sa == sb:
                continue
            n_mismatch += 1
            only_a, only_b = sorted(sa - sb), sorted(sb - sa)
            q = query[b, m].astype(np.float64)
            da = np.sum((xyz[b, only_a].astype(np.float64) - q) ** 2, -1)
            db = np.sum((xyz[b, only_b].astype(np.float64) - q) ** 2, -1)
            # every point we picked instead must be at least as near (f64) as every point the reference picked instead
            assert da.max() <= db.min() * (1 + 1e-6) + 1e-12, \
                f"{what}: query ({b},{m}) ours-only {only_a} d2={da} vs ref-only {only_b} d2={db}"
    return n_mismatch
