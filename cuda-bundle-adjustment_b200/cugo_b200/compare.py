"""Comparison report of two optimisation runs of the same graph: the per-iteration chi2 table and the RMSE of
the final estimates, as the reference's comparison sample prints them
(samples/sample_comparison_with_g2o/main.cpp:101-148: "iteration | chi2 CPU | chi2 GPU", then
"Rotation / Translation / Landmark" RMSE over quaternion coefficients, translations and landmark positions).
That sample compares against g2o, which cannot be built here (no Eigen, no network); the second run is whatever the
caller supplies -- the CPU oracle, the reference GPU library's dump, or another engine configuration.
"""
from __future__ import annotations

import numpy as np


def rmse_report(poses_a, lms_a, poses_b, lms_b):
    """poses [P,7] (qx qy qz qw tx ty tz), landmarks [L,3] -> dict(rotation, translation, landmark).
    Quaternion coefficients are compared as stored (both sides normalise with w >= 0,
    src/cuda/cuda_block_solver.cu:764-775)."""
    pa, pb = np.asarray(poses_a, np.float64), np.asarray(poses_b, np.float64)
    la, lb = np.asarray(lms_a, np.float64), np.asarray(lms_b, np.float64)
    nP, nL = max(pa.shape[0], 1), max(la.shape[0], 1)
    return dict(rotation=float(np.sqrt(((pa[:, :4] - pb[:, :4]) ** 2).sum() / nP)),
                translation=float(np.sqrt(((pa[:, 4:] - pb[:, 4:]) ** 2).sum() / nP)),
                landmark=float(np.sqrt(((la - lb) ** 2).sum() / nL)))


def chi2_table(chi2_a, chi2_b, name_a="A", name_b="B"):
    """The reference's table, one line per iteration; a run that stopped early shows N/A."""
    n = max(len(chi2_a), len(chi2_b))
    lines = ["%10s|%16s|%16s|%10s" % ("iteration", "chi2 " + name_a, "chi2 " + name_b, "rel.diff")]
    for i in range(n):
        a = "%16.6f" % chi2_a[i] if i < len(chi2_a) else "%16s" % "N/A"
        b = "%16.6f" % chi2_b[i] if i < len(chi2_b) else "%16s" % "N/A"
        r = ("%10.2e" % (abs(chi2_a[i] - chi2_b[i]) / max(abs(chi2_b[i]), 1e-300))
             if i < len(chi2_a) and i < len(chi2_b) else "%10s" % "N/A")
        lines.append("%10d|%s|%s|%s" % (i + 1, a, b, r))
    return "\n".join(lines)


def report(chi2_a, est_a, chi2_b, est_b, name_a="A", name_b="B"):
    r = rmse_report(est_a[0], est_a[1], est_b[0], est_b[1])
    return (chi2_table(chi2_a, chi2_b, name_a, name_b) +
            "\n=== RMSE between %s estimates and %s estimates :\nRotation    : %.2e\nTranslation : %.2e\nLandmark    : %.2e" %
            (name_a, name_b, r["rotation"], r["translation"], r["landmark"])), r
