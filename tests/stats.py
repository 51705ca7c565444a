"""Chi-square helpers for the sampling-law tests (TEST INFRASTRUCTURE)."""
from __future__ import annotations

import numpy as np
from scipy.stats import chi2


def two_sample_chi2(a, b, min_expected=5.0):
    """Homogeneity test of two count vectors over the same categories (different totals allowed).
    Categories whose expected count in the SMALLER sample is below `min_expected` are pooled into one (the
    chi-square approximation needs that; it matters when a 1 000-draw sample meets a 400 000-draw simulation).
    -> (statistic, dof, p)"""
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    assert a.shape == b.shape
    tot = a.sum() + b.sum()
    small = (a + b) * min(a.sum(), b.sum()) / max(tot, 1.0) < min_expected
    if small.any() and (~small).any():
        a = np.append(a[~small], a[small].sum())
        b = np.append(b[~small], b[small].sum())
    keep = (a + b) > 0
    a, b = a[keep], b[keep]
    A, B = a.sum(), b.sum()
    if len(a) < 2 or A == 0 or B == 0:
        return 0.0, 0, 1.0
    stat = float((((a * np.sqrt(B / A) - b * np.sqrt(A / B)) ** 2) / (a + b)).sum())
    dof = len(a) - 1
    return stat, dof, float(chi2.sf(stat, dof))


def one_sample_chi2(obs, expected, min_expected=5.0):
    """Goodness of fit of counts against expected counts (same total).  -> (statistic, dof, p)"""
    obs, expected = np.asarray(obs, np.float64).ravel(), np.asarray(expected, np.float64).ravel()
    small = expected < min_expected
    if small.any() and (~small).any():
        obs = np.append(obs[~small], obs[small].sum())
        expected = np.append(expected[~small], expected[small].sum())
    keep = expected > 0
    obs, expected = obs[keep], expected[keep]
    stat = float((((obs - expected) ** 2) / expected).sum())
    dof = len(obs) - 1
    return stat, dof, float(chi2.sf(stat, dof)) if dof > 0 else 1.0
