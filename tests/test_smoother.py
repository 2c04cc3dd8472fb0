"""cg_eigs + Chebyshev smoother against the outputs of the unmodified reference
(tests/golden/*: cheby_* records written by oracle/ref_driver.c --cheby)."""
import numpy as np
import pytest

from oracle_lib import Oracle
from solver_ref import cg_eigs, cheby

CASES = ["brick_l2_p3", "sphere_p2", "sphere_hp_amr1_zero"]


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def problem(A):
    imax, eigs_imax, ratio, lmax = (float(x) for x in A["cheby_params"])
    return int(imax), int(eigs_imax), ratio, lmax, np.ascontiguousarray(A["cheby_rhs"]), 0.5 * A["u"]


@pytest.mark.parametrize("case", CASES)
def test_restatement_matches_reference(golden, case):
    """pins tests/solver_ref.py (and the oracle apply under it) to the reference"""
    A = golden(case)
    imax, eigs_imax, ratio, lmax, rhs, u0 = problem(A)
    orc = Oracle(A)
    e_new, _ = cg_eigs(orc.apply, u0, rhs, eigs_imax, use_new=1)
    e_old, u1 = cg_eigs(orc.apply, u0, rhs, eigs_imax, use_new=0)
    assert abs(e_new - A["cheby_eig_new"][0]) <= 1e-10 * e_new
    assert abs(e_old - A["cheby_eig_old"][0]) <= 1e-10 * e_old
    assert e_old == pytest.approx(lmax, rel=1e-10)
    assert rel(u1, A["cheby_u_after_eigs"]) < 1e-10
    u2, r2 = cheby(orc.apply, A["cheby_u_after_eigs"], rhs, imax, lmax / ratio, lmax)
    assert rel(u2, A["cheby_u"]) < 1e-11
    assert rel(r2, A["cheby_r"]) < 1e-10


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_gpu_cg_eigs_and_cheby_match_reference(golden, case):
    from disco4est_b200 import Mesh

    A = golden(case)
    imax, eigs_imax, ratio, lmax, rhs, u0 = problem(A)
    m = Mesh(A)
    u = u0.copy()
    e_new = m.cg_eigs(u, rhs, eigs_imax, use_new=1)
    assert abs(e_new - A["cheby_eig_new"][0]) <= 1e-10 * e_new
    u = u0.copy()
    e_old = m.cg_eigs(u, rhs, eigs_imax, use_new=0)
    assert abs(e_old - A["cheby_eig_old"][0]) <= 1e-10 * e_old
    assert rel(u, A["cheby_u_after_eigs"]) < 1e-10       # the estimate advances u in place, like the reference
    # zero-guess variant leaves u alone
    u_keep = u0.copy()
    m.cg_eigs(u_keep, rhs, eigs_imax, use_new=0, use_zero_guess=1)
    assert np.array_equal(u_keep, u0)
    u = np.ascontiguousarray(A["cheby_u_after_eigs"]).copy()
    r = np.zeros_like(u)
    m.cheby_smooth(u, rhs, r, imax, e_old / ratio, e_old)
    assert rel(u, A["cheby_u"]) < 1e-11
    assert rel(r, A["cheby_r"]) < 1e-10
    m.close()


@pytest.mark.gpu
def test_gpu_cheby_device_pointers(golden):
    import torch
    from disco4est_b200 import Mesh

    A = golden("brick_l2_p3")
    imax, eigs_imax, ratio, lmax, rhs, u0 = problem(A)
    m = Mesh(A)
    u = torch.from_numpy(np.ascontiguousarray(A["cheby_u_after_eigs"])).cuda()
    r = torch.zeros_like(u)
    m.cheby_smooth(u, torch.from_numpy(rhs).cuda(), r, imax, lmax / ratio, lmax)
    m.sync()
    assert rel(u.cpu().numpy(), A["cheby_u"]) < 1e-11
    assert rel(r.cpu().numpy(), A["cheby_r"]) < 1e-10
    m.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["sphere_p_amr1", "sphere_p_amr1_robin"])
@pytest.mark.parametrize("split", ["0", "1"])
def test_fused_smoother_on_hp_meshes(golden, case, split, monkeypatch):
    """meshes whose irregular elements run in the mortar kernel (element-wise or face split) take the fused
    Chebyshev epilogue like the conforming ones: same iterate and residual as the unfused sequence and as the
    restatement on top of the oracle"""
    from disco4est_b200 import Mesh

    A = golden(case)
    orc = Oracle(A)
    rng = np.random.default_rng(8)
    rhs, u0 = rng.standard_normal(orc.n), rng.standard_normal(orc.n)
    lmax, _ = cg_eigs(orc.apply, u0.copy(), rhs, 6, use_new=0)
    u_ref, r_ref = cheby(orc.apply, u0.copy(), rhs, 5, lmax / 30.0, lmax)
    out = {}
    monkeypatch.setenv("D4B_SPLIT", split)
    for fused in ("0", "1"):
        monkeypatch.setenv("D4B_NO_FUSED_CHEBY", "0" if fused == "1" else "1")
        m = Mesh(A)
        assert m.hp_elements() > 0
        u, r = u0.copy(), np.zeros_like(u0)
        m.cheby_smooth(u, rhs, r, 5, lmax / 30.0, lmax)
        m.close()
        out[fused] = (u, r)
        assert rel(u, u_ref) < 1e-11 and rel(r, r_ref) < 1e-10
    assert rel(out["1"][0], out["0"][0]) < 1e-13


@pytest.mark.gpu
@pytest.mark.parametrize("p", [1, 2, 3, 4, 5, 6, 7])
def test_fused_solvers_on_bricks(p, monkeypatch):
    """meshes whose elements all run in the tangential-free kernel: cg_eigs with d.Ad taken from the apply
    kernel (energy form) and the direction update riding on the interpolation kernel, and the Chebyshev
    recurrence carried in the Gauss basis (one kernel per sweep) -- against the unfused sequence of apply +
    vector kernels (D4B_FUSED_SOLVERS=0) and the restatement on top of the oracle"""
    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_desc

    A = brick_desc(1 if p > 4 else 2, p)
    orc = Oracle(A)
    rng = np.random.default_rng(80 + p)
    rhs, u0 = rng.standard_normal(orc.n), rng.standard_normal(orc.n)
    lmax_ref, u1_ref = cg_eigs(orc.apply, u0.copy(), rhs, 8, use_new=0)
    u_ref, r_ref = cheby(orc.apply, u0.copy(), rhs, 6, lmax_ref / 30.0, lmax_ref)
    out = {}
    for fused in ("0", "1"):
        monkeypatch.setenv("D4B_FUSED_SOLVERS", fused)
        m = Mesh(A)
        assert m.flat_elements() == int(A["n_local"])
        u = u0.copy()
        lmax = m.cg_eigs(u, rhs, 8, use_new=0)
        assert abs(lmax - lmax_ref) <= 1e-10 * lmax_ref
        assert rel(u, u1_ref) < 1e-10
        us, r = u0.copy(), np.zeros_like(u0)
        m.cheby_smooth(us, rhs, r, 6, lmax_ref / 30.0, lmax_ref)
        assert rel(us, u_ref) < 1e-11 and rel(r, r_ref) < 1e-10
        m.close()
        out[fused] = (lmax, u, us, r)
    assert abs(out["1"][0] - out["0"][0]) <= 1e-12 * out["0"][0]
    assert rel(out["1"][1], out["0"][1]) < 1e-12 and rel(out["1"][2], out["0"][2]) < 1e-12
    assert rel(out["1"][3], out["0"][3]) < 1e-11
