"""Python restatement of the reference's lambda_max estimate and Chebyshev
smoother over an arbitrary `apply(u) -> Au` callable.  TEST INFRASTRUCTURE.

cg_eigs:  src/Solver/d4est_solver_cg_eigs.c:115-276 (Gershgorin bounds :9-65)
cheby:    src/Solver/d4est_solver_multigrid_smoother_cheby.c:81-176
"""
import math

import numpy as np


def gershgorin(use_new, i, n, a0, b0, a1, b1):
    """tridiag_gershgorin (:9-36) / tridiag_gershgorin_new (:38-65): upper bound only."""
    if not use_new:
        if i != 0 and i < n - 1:
            diag = 1.0 / a1 + b0 / a0
            off = abs(math.sqrt(b1) / a1) + abs(math.sqrt(b0) / a0)
        elif i == 0:
            diag = 1.0 / a1
            off = math.sqrt(b1) / a1
        else:
            diag = 1.0 / a1 + b0 / a0
            off = abs(math.sqrt(b0) / a0)
    else:
        if i != 0:
            diag = 1.0 / a1 + b0 / a0
            off = abs(math.sqrt(b0) / a0)
        else:
            diag = 1.0 / a1
            off = math.sqrt(b1) / a1
    return diag + off


def cg_eigs(apply, u, rhs, imax, use_new=0):
    """Returns (spectral_bound, u advanced by imax CG steps) -- the reference
    updates vecs->u in place (:229)."""
    u = u.copy()
    n = u.size
    r = rhs - apply(u)
    d = r.copy()
    delta_new = float(r @ r)
    alpha = beta = -1.0
    bound = 0.0
    for i in range(imax):
        Ad = apply(d)
        alpha_old = alpha
        alpha = delta_new / float(d @ Ad)
        u += alpha * d
        r -= alpha * Ad
        delta_old = delta_new
        delta_new = float(r @ r)
        beta_old = beta
        beta = delta_new / delta_old
        d = r + beta * d
        t = gershgorin(use_new, i, n, alpha_old, beta_old, alpha, beta)
        bound = max(bound, t) if i > 0 else t
    return bound, u


def cheby(apply, u, rhs, iters, lmin, lmax):
    """Returns (u, r = rhs - A u) after `iters` Chebyshev iterations."""
    u = u.copy()
    d = 0.5 * (lmax + lmin)
    c = 0.5 * (lmax - lmin)
    p = np.zeros_like(u)
    alpha = 0.0
    for i in range(iters):
        r = rhs - apply(u)
        if i == 0:
            alpha = 1.0 / d
        elif i == 1:
            alpha = 2.0 * d / (2.0 * d * d - c * c)
        else:
            alpha = 1.0 / (d - alpha * c * c / 4.0)
        beta = alpha * d - 1.0
        p = alpha * r + beta * p
        u += p
    return u, rhs - apply(u)
