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


# ---- multigrid -------------------------------------------------------------------
_LGL = {}


def lgl_nodes(N):
    if N not in _LGL:
        from disco4est_b200.brick import lgl_nodes as f
        _LGL[N] = f(N)
    return _LGL[N]


_P1D = {}


def prolong_1d(degH, degh, half):
    key = (degH, degh, half)
    if key not in _P1D:
        _P1D[key] = _prolong_1d(degH, degh, half)
    return _P1D[key]


def _prolong_1d(degH, degh, half):
    """[nh x nH] Lagrange basis on LGL(degH) at the LGL(degh) nodes of the whole
    interval (half 0: build_p_prolong_1d, d4est_operators.c:995-1012) or of its lower /
    upper half (half 1, 2: build_hp_prolong_1d :944-993)."""
    xH, xh = lgl_nodes(degH + 1), lgl_nodes(degh + 1)
    y = xh if half == 0 else (0.5 * (xh - 1.0) if half == 1 else 0.5 * (xh + 1.0))
    if half == 0 and degH == degh:
        return np.eye(degh + 1)
    P = np.ones((degh + 1, degH + 1))
    for j in range(degH + 1):
        for k in range(degH + 1):
            if k != j:
                P[:, j] *= (y - xH[k]) / (xH[j] - xH[k])
    return P


def prolong_3d(degH, degh, child, count):
    h = (0, 0, 0) if count == 1 else (1 + (child & 1), 1 + ((child >> 1) & 1), 1 + ((child >> 2) & 1))
    return np.kron(prolong_1d(degH, degh, h[2]), np.kron(prolong_1d(degH, degh, h[1]), prolong_1d(degH, degh, h[0])))


class Transfer:
    """coarse level <-> fine level, (child_first, child_count) per coarse element
    (d4est_solver_multigrid_callbacks.h:110-203 restriction, :237-329 prolongation)."""

    def __init__(self, coarse, fine, child_first, child_count):
        self.c, self.f, self.first, self.count = coarse, fine, child_first, child_count

    def _blocks(self):
        for e in range(int(self.c["n_local"])):
            dH = int(self.c["deg"][e]); sH = int(self.c["nodal_stride"][e])
            for ch in range(int(self.count[e])):
                ef = int(self.first[e]) + ch
                dh = int(self.f["deg"][ef]); sh = int(self.f["nodal_stride"][ef])
                yield sH, (dH + 1) ** 3, sh, (dh + 1) ** 3, prolong_3d(dH, dh, ch, int(self.count[e]))

    def restrict(self, fine):
        out = np.zeros(int(self.c["local_nodes"]))
        for sH, nH, sh, nh, P in self._blocks():
            out[sH:sH + nH] += P.T @ fine[sh:sh + nh]
        return out

    def prolong(self, coarse):
        out = np.zeros(int(self.f["local_nodes"]))
        for sH, nH, sh, nh, P in self._blocks():
            out[sh:sh + nh] = P @ coarse[sH:sH + nH]
        return out


def bottom_cg(apply, rhs, imax, rtol, atol):
    """d4est_solver_multigrid_bottom_solver_cg.c:48-200, zero initial guess"""
    u = np.zeros_like(rhs)
    r = rhs - apply(u)
    d = r.copy()
    delta_new = float(r @ r)
    delta_0 = delta_new
    i = 0
    while i < imax and delta_new > atol * atol + delta_0 * rtol * rtol:
        Ad = apply(d)
        alpha = delta_new / float(d @ Ad)
        u += alpha * d
        r -= alpha * Ad
        delta_old = delta_new
        delta_new = float(r @ r)
        d = r + (delta_new / delta_old) * d
        i += 1
    return u, i


def vcycle(applies, transfers, u, rhs, cheby_imax, eigs_imax, ratio, mult=1.0, use_new=0,
           bottom=(100, 1e-10, 0.0), bottom_cheby=None):
    """One V-cycle, d4est_solver_multigrid.c:751-1349 (eigenvalues recomputed on both legs).
    applies[l], transfers[l] (between l and l+1), coarsest first.  Returns (u, r2, eigs).
    bottom = (imax, rtol, atol) of mg_bottom_solver_cg, or bottom_cheby = (cheby_imax, eigs_cg_imax, ratio)
    of mg_bottom_solver_cheby (d4est_solver_multigrid_bottom_solver_cheby.c:60-115)."""
    top = len(applies) - 1
    us, rhss, eigs = [None] * (top + 1), [None] * (top + 1), [0.0] * (top + 1)
    us[top], rhss[top] = u.copy(), rhs
    r = None

    def smooth(l):
        e, us[l] = cg_eigs(applies[l], us[l], rhss[l], eigs_imax, use_new)
        eigs[l] = e * mult
        us[l], rr = cheby(applies[l], us[l], rhss[l], cheby_imax, eigs[l] / ratio, eigs[l])
        return rr

    for l in range(top, 0, -1):
        if l != top:
            us[l] = np.zeros_like(rhss[l])
        r = smooth(l)
        rhss[l - 1] = transfers[l - 1].restrict(r)
    if bottom_cheby is None:
        us[0], _ = bottom_cg(applies[0], rhss[0], *bottom)
    else:
        e0, u0 = cg_eigs(applies[0], np.zeros_like(rhss[0]), rhss[0], bottom_cheby[1], use_new)
        us[0], _ = cheby(applies[0], u0, rhss[0], bottom_cheby[0], e0 / bottom_cheby[2], e0)
    for l in range(0, top):
        us[l + 1] = us[l + 1] + transfers[l].prolong(us[l])
        r = smooth(l + 1)
    return us[top], float(r @ r), eigs


# ---- outer Krylov ----------------------------------------------------------------
def cg(apply, u, rhs, imax, rtol, atol=0.0):
    """d4est_solver_cg_solve (src/Solver/d4est_solver_cg.c:75-197); returns (u, iterations)"""
    u = u.copy()
    r = rhs - apply(u)
    d = r.copy()
    delta_new = float(r @ r)
    delta_0 = delta_new
    its = 0
    while its < imax and delta_new > atol * atol + delta_0 * rtol * rtol:
        Ad = apply(d)
        alpha = delta_new / float(d @ Ad)
        u += alpha * d
        r -= alpha * Ad
        delta_old = delta_new
        delta_new = float(r @ r)
        d = r + (delta_new / delta_old) * d
        its += 1
    return u, its


def fcg(apply, pc, u, rhs, imax, rtol, atol=0.0, textbook=False):
    """d4est_solver_fcg_solve (src/Solver/d4est_solver_fcg_improved.c:103-346): flexible CG, truncation 1.
    pc(r) -> v or None for the identity.  Returns (u, iterations, [|r_k|^2 at the start of iteration k]).

    What the reference EXECUTES differs from the recurrence its comments state: the direction updates
    are written d4est_linalg_vec_axpyeqz(c, d, v, d) (:260-262), and axpyeqz copies y into z before
    adding c x (src/LinearAlgebra/d4est_linalg.c:217-221) -- with z aliasing x that yields
    d_k = (1 + c) v_k and q_k = (1 + c) w_k, not v_k + c d_{k-1}.  Parity means reproducing that
    (textbook=False); textbook=True is the recurrence of the comments (Notay & Napov)."""
    u = u.copy()
    r = rhs - apply(u)
    tol = atol + rtol * np.sqrt(float(r @ r))
    hist = []
    rho = 0.0
    d = q = None
    for k in range(imax):
        hist.append(float(r @ r))
        v = pc(r) if pc is not None else r.copy()
        w = apply(v)
        alpha, beta = float(v @ r), float(v @ w)
        if k > 0:
            gamma, rr = float(v @ q), float(r @ r)
            c = -gamma / rho
            d = v + c * (d if textbook else v)
            q = w + c * (q if textbook else w)
            rho = beta - gamma * gamma / rho
        else:
            rho, d, q = beta, v.copy(), w.copy()
        u += (alpha / rho) * d
        r -= (alpha / rho) * q
        if k > 0 and np.sqrt(rr) <= tol:
            break
    return u, len(hist), hist


# ---- Newton-Krylov-multigrid on a mass-type nonlinearity -----------------------------
def vcycle2(ap1, ap0, transfer, rhs, cheby_imax, eigs_imax, ratio, use_new=0, bottom=(15, 15, 30.0)):
    """One two-level V-cycle from a zero guess with a Chebyshev bottom solver
    (d4est_solver_multigrid_bottom_solver_cheby.c:60-115): what d4est_krylov_pc_multigrid_apply does with
    vcycle_imax = 1.  Returns u."""
    def smooth(ap, u, b, imax, eimax, rat):
        e, u = cg_eigs(ap, u, b, eimax, use_new)
        return cheby(ap, u, b, imax, e / rat, e)

    u, r = smooth(ap1, np.zeros_like(rhs), rhs, cheby_imax, eigs_imax, ratio)
    rc = transfer.restrict(r)
    uc, _ = smooth(ap0, np.zeros_like(rc), rc, bottom[0], bottom[1], bottom[2])
    u, r = smooth(ap1, u + transfer.prolong(uc), rhs, cheby_imax, eigs_imax, ratio)
    return u


def newton(residual, linearize, u, atol, rtol, imin, imax, fcg_rtol, fcg_atol, fcg_imax):
    """d4est_solver_newton_solve (src/Solver/d4est_solver_newton.c:136-367) with d4est_solver_fcg_solve as the
    Krylov callback.  residual(u) -> F; linearize(u0) -> (apply, pc).  Returns (u, FCG iterations per step,
    residual norms)."""
    x = u.copy()
    f0 = residual(x)
    fnrm = float(np.sqrt(f0 @ f0))
    norms, its = [fnrm], []
    stop = atol + rtol * fnrm
    itc = 0
    while (fnrm > stop or itc < imin) and itc < imax:
        ap, pc = linearize(x)
        step, n, _ = fcg(ap, pc, np.zeros_like(x), -f0, fcg_imax, fcg_rtol, fcg_atol)
        its.append(n)
        x = x + step
        f0 = residual(x)
        fnrm = float(np.sqrt(f0 @ f0))
        norms.append(fnrm)
        itc += 1
    return x, its, norms
