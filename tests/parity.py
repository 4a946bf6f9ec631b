"""Comparison rules of the parity tests (tolerances of BASELINE.json's north_star)."""
import numpy as np

LAMBDA_RTOL = 1e-4      # |Δλ| <= 1e-4 * max(|λ|, LAMBDA_FLOOR * λmax)
# Floor: the reference solver (Eigen fp32 tridiagonal QR) carries an absolute error of a few
# fp32 ulps of λmax on EVERY eigenvalue, so a small λ cannot be reproduced to 1e-4 of itself by
# anyone (SURVEY §7 hard parts).  1e-4 * 1e-2 * λmax = 1e-6 λmax ≈ 8 ulps of λmax.  The GPU path
# solves the same fp32 moment in fp64; test_gpu_pca.py::test_eigen_closer_to_fp64_truth shows it is
# the oracle, not the GPU, that sits this far from the exact eigenvalues.
LAMBDA_FLOOR = 1e-2
NORMAL_TOL = 1e-4       # 1 - |n · n_ref| <= 1e-4 where the spectral gap allows
GAP_MIN = 1e-3          # eigenvector compared only if its eigenvalue is separated by >= 1e-3 * λmax
FEATURE_RTOL = 3e-4


def check_features(got, ref, valid=None, what=""):
    """got/ref: FEATURE_DTYPE arrays.  Index/int fields exact; λ, vectors, features within tolerance."""
    assert got.shape == ref.shape, what
    np.testing.assert_array_equal(got["ptID"], ref["ptID"], err_msg=what + " ptID")
    np.testing.assert_array_equal(got["ptNum"], ref["ptNum"], err_msg=what + " ptNum")
    m = np.ones(got.shape[0], bool) if valid is None else valid.astype(bool)
    lg, lr = got["la"][m].astype(np.float64), ref["la"][m].astype(np.float64)
    lmax = np.abs(lr).max(axis=1, keepdims=True)
    tol = LAMBDA_RTOL * np.maximum(np.abs(lr), LAMBDA_FLOOR * lmax) + 1e-30
    bad = np.abs(lg - lr) > tol
    assert not bad.any(), f"{what}: {bad.sum()} eigenvalues off; worst {np.abs(lg - lr)[bad].max():.3e}"
    # eigenvectors up to sign, gated on the spectral gap
    gaps = np.stack([lr[:, 1] - lr[:, 0], np.minimum(lr[:, 1] - lr[:, 0], lr[:, 2] - lr[:, 1]), lr[:, 2] - lr[:, 1]], 1)
    stats = {}
    for k, name in enumerate(("v1", "v2", "v3")):
        vg, vr = got[name][m].astype(np.float64), ref[name][m].astype(np.float64)
        ok = gaps[:, k] >= GAP_MIN * lmax[:, 0]
        dots = np.abs((vg * vr).sum(1))
        err = 1.0 - dots[ok]
        stats[name] = (int(ok.sum()), float(err.max()) if err.size else 0.0)
        assert not (err > NORMAL_TOL).any(), f"{what}: {name} off by {err.max():.3e} on {int((err > NORMAL_TOL).sum())} records"
        nrm = np.linalg.norm(vg, axis=1)
        assert np.allclose(nrm, 1.0, atol=1e-5), what + " unit norm"
    # structure features: only where la1 is resolved (else sqrt of round-off)
    okf = (lr[:, 0] > LAMBDA_FLOOR * lmax[:, 0]) & (lg[:, 0] > 0)
    for name in ("lineStruc", "planeStruc", "scatterStruc"):
        fg, fr = got[name][m][okf].astype(np.float64), ref[name][m][okf].astype(np.float64)
        bad = np.abs(fg - fr) > FEATURE_RTOL * np.maximum(1.0, np.abs(fr))
        assert not bad.any(), f"{what}: {name} off on {bad.sum()} records"
    return stats


def check_stale_equal(got, ref, invalid, what=""):
    """Carry-forward records must repeat a valid record's payload: compare them with the same rules."""
    return check_features(got[invalid], ref[invalid], None, what + " (stale)") if invalid.any() else {}
