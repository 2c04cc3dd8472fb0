"""Volume quadrature operators and the fused mass-type diagonal.

Reference side: oracle/ref_driver.c --quad calls, element by element,
d4est_quadrature_interpolate / _apply_galerkin_integral / _apply_mass_matrix /
_apply_fofufofvlilj / _apply_fofufofvlj (src/Quadrature/d4est_quadrature.c) of the
unmodified reference with a ConstantDensityStar-like f; the outputs are in the fixtures."""
import numpy as np
import pytest

from oracle_lib import Oracle

CASES = ["brick_l2_p3", "sphere_p2", "sphere_hp_amr1_zero"]


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("case", CASES)
def test_oracle_matches_reference(golden, case):
    A = golden(case)
    O = Oracle(A)
    u = np.ascontiguousarray(A["u"])
    J = np.ascontiguousarray(A["J_quad"])
    assert rel(O.interp_to_quad(u), A["q_interp"]) < 1e-14
    assert rel(O.apply_galerkin(A["q_interp"], J), A["q_galerkin"]) < 1e-14
    assert rel(O.apply_mass(u, J), A["q_mass"]) < 1e-14
    # Newton-linearised mass-type term: V^T diag(w J f(x, V u0)) V u
    assert rel(O.apply_mass(u, A["q_jf"]), A["q_jacterm"]) < 1e-14
    # mass matrix == Galerkin integral of the interpolant (d4est_quadrature.c:385-477)
    assert rel(A["q_mass"], A["q_galerkin"]) < 1e-14


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_gpu_quadrature_ops_match_reference(golden, case):
    from disco4est_b200 import Mesh

    A = golden(case)
    m = Mesh(A)
    u = np.ascontiguousarray(A["u"])
    J = np.ascontiguousarray(A["J_quad"])
    uq = np.zeros(A["q_interp"].size)
    m.interp_to_quad(u, uq)
    assert rel(uq, A["q_interp"]) < 1e-13
    out = np.zeros(u.size)
    m.apply_galerkin(np.ascontiguousarray(A["q_interp"]), J, out)
    assert rel(out, A["q_galerkin"]) < 1e-13
    m.apply_mass(u, J, out)
    assert rel(out, A["q_mass"]) < 1e-13
    m.apply_mass(u, np.ascontiguousarray(A["q_jf"]), out)
    assert rel(out, A["q_jacterm"]) < 1e-13
    # residual-type term of apply_fofufofvlj: V^T (w J f(x, V u0)); f/J evaluated by the caller.
    # q_jf = J f_jac(u0q); the residual functor is f_res = f_jac (1 + u0q) / 5
    u0q = np.zeros(uq.size)
    m.interp_to_quad(np.ascontiguousarray(A["q_u0"]), u0q)
    m.apply_galerkin(np.ascontiguousarray(A["q_jf"] * (1.0 + u0q) / 5.0), None, out)
    assert rel(out, A["q_resterm"]) < 1e-12
    # device pointers give the same bits
    import torch

    du = torch.from_numpy(u).cuda()
    dj = torch.from_numpy(np.ascontiguousarray(A["q_jf"])).cuda()
    dout = torch.zeros(u.size, dtype=torch.float64, device="cuda")
    m.apply_mass(du, dj, dout)
    m.sync()
    assert np.array_equal(dout.cpu().numpy(), _mass(m, u, A["q_jf"]))
    m.close()


def _mass(m, u, jf):
    out = np.zeros(u.size)
    m.apply_mass(u, np.ascontiguousarray(jf), out)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("force_generic", [False, True])
def test_gpu_apply_with_fused_mass_diagonal(golden, case, force_generic, monkeypatch):
    """Jacobian apply of the nonlinear problems: SIPG Laplacian + fofufofvlilj term in one pass
    (constant_density_star_apply_jac, constant_density_star_fcns.h:528-600)"""
    from disco4est_b200 import Mesh

    if force_generic:
        monkeypatch.setenv("D4B_FORCE_GENERIC", "1")
    A = golden(case)
    m = Mesh(A)
    u = np.ascontiguousarray(A["u"])
    want = A["Au_plain"] + A["q_jacterm"]
    Au = np.zeros(u.size)
    m.set_mass_diag(np.ascontiguousarray(A["q_jf"]))
    m.apply(u, Au)
    assert rel(Au, want) < 1e-12
    # stiffness-only entry is unaffected, removing the diagonal restores the Laplacian
    m.set_mass_diag(None)
    m.apply(u, Au)
    assert rel(Au, A["Au_plain"]) < 1e-12
    m.close()


@pytest.mark.gpu
def test_gpu_smoother_with_fused_mass_diagonal(golden):
    """the fused Chebyshev epilogue on the Jacobian operator == apply + vector update"""
    from disco4est_b200 import Mesh
    from solver_ref import cheby

    A = golden("brick_l2_p3")
    O = Oracle(A)
    jf = -np.ascontiguousarray(A["q_jf"])          # positive shift keeps the operator SPD
    m = Mesh(A)
    m.set_mass_diag(jf)
    u = np.ascontiguousarray(A["u"])
    op = lambda x: O.apply(x) + O.apply_mass(x, jf)
    rhs = op(u)
    lmax = m.cg_eigs(u.copy() * 0.5, rhs, 10)
    u1 = u * 0.5
    r1 = np.zeros(u.size)
    m.cheby_smooth(u1, rhs, r1, 8, lmax / 30.0, lmax)
    u2, r2 = cheby(op, u * 0.5, rhs, 8, lmax / 30.0, lmax)
    assert rel(u1, u2) < 1e-11 and rel(r1, r2) < 1e-9
    m.close()
