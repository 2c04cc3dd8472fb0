"""BASELINE.json configs[4] in small: the two-puncture problem on the 13-tree cubed sphere with a Robin outer
boundary, solved by the reference's in-house Newton -> FCG -> multigrid stack
(src/Problems/TwoPunctures/multi_punctures_anares_inhouse.c:849-964).  The fixture was written by the
UNMODIFIED reference (oracle/ref_driver.c --newton, oracle/make_golden.sh): residual and Jacobian samples
evaluated with its own puncture callbacks, Newton steps, FCG iterations per step, residual norms, final u.

Device side: the puncture terms are device functors (d4b_mesh_set_nonlinear_term), the residual and the
linearisation run on the device, the coarse level applies the reference's Galerkin mass term P^T M P
(d4b_mg_set_coarse_mass) and d4b_newton_solve keeps every vector in HBM."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-12


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def _robin(A, R2):
    """Robin data of multi_punctures_robin_coeff_sphere_fcn / _robin_bc_rhs_fcn: coefficient 1 / r = 1 / R2 on the
    outer sphere, right-hand side 0 (only boundary sides read these arrays)"""
    mp = int(np.asarray(A["local_mortar_nodes_quad"]).reshape(-1)[0])
    A = dict(A)
    A["bc_type"] = 1
    A["robin_coeff_mortar_quad"] = np.full(mp, 1.0 / R2)
    A["robin_rhs_mortar_quad"] = np.zeros(mp)
    return A


@pytest.fixture(scope="module")
def problem(golden):
    from disco4est_b200 import Mesh

    A = golden("sphere_punctures_newton")
    m = Mesh(A, device=0)
    m.set_nonlinear_term("punctures", punctures=A["punctures"], xyz_quad=A["xyz_quad"], jac_quad=A["J_quad"])
    yield A, m
    m.close()


def test_fixture_is_the_reference_run(golden):
    A = golden("sphere_punctures_newton")
    assert A["nw_fcg_iterations"].astype(int).tolist() == [3, 3, 8, 2, 4, 2]
    assert len(A["nw_residual_norms"]) == 7 and A["nw_residual_norms"][-1] < 1e-8 * A["nw_residual_norms"][0]


def test_residual_matches_reference(problem):
    """F(u) = A u + V^T w J f(x, V u), f = -K^2/8 psi^-7 evaluated on the device, against
    multi_punctures_build_residual with the reference's own callbacks"""
    A, m = problem
    F = np.zeros_like(A["nw_us"])
    m.build_residual(np.ascontiguousarray(A["nw_us"]), F)
    assert rel(F, A["nw_F_us"]) < TOL


def test_jacobian_apply_matches_reference(problem):
    """J(u0) v = A v + V^T diag(w J 7/8 K^2 psi(u0)^-8) V v against multi_punctures_apply_jac"""
    A, m = problem
    m.linearize(np.ascontiguousarray(A["nw_us"]))
    Jv = np.zeros_like(A["nw_v"])
    m.apply(np.ascontiguousarray(A["nw_v"]), Jv)
    m.linearize(None)
    assert rel(Jv, A["nw_J_us_v"]) < TOL


def test_nonlinear_term_with_device_geometry(problem):
    """the same residual on a mesh whose factors AND quadrature-point positions are generated on the device"""
    from disco4est_b200 import Mesh
    from disco4est_b200.sphere import sphere_desc

    A, _ = problem
    T, g = sphere_desc(1, p=2, prefactor=2.0, radii=(1.0, 2.0, 3.0))
    m = Mesh(_robin(T, 3.0), device=0, geom=g)
    m.set_nonlinear_term("punctures", punctures=A["punctures"])
    F = np.zeros_like(A["nw_us"])
    m.build_residual(np.ascontiguousarray(A["nw_us"]), F)
    m.close()
    assert rel(F, A["nw_F_us"]) < 1e-11


def test_coarse_level_galerkin_mass_term(problem):
    """mode "galerkin": level 0 of the hierarchy applies A_0 + P^T M_1 P (what the reference's multigrid matrix
    operator is documented to do): checked against prolong -> fine mass -> restrict done by hand"""
    import torch

    from disco4est_b200 import Mesh
    from disco4est_b200.runtime import Multigrid, cheby_params
    from disco4est_b200.sphere import sphere_desc

    A, m1 = problem
    T0, g0 = sphere_desc(0, p=2, prefactor=2.0, radii=(1.0, 2.0, 3.0))
    m0 = Mesh(_robin(T0, 3.0), device=0, geom=g0)
    tr = [(8 * np.arange(13, dtype=np.int32), np.full(13, 8, np.int32))]
    mg = Multigrid([m0, m1], tr, cheby_params(8, 10, 30.0, 1.0), bottom_solver="cheby", bottom_cheby=cheby_params(15, 15, 30.0, 1.0))
    rng = np.random.default_rng(3)
    uc = rng.standard_normal(m0.n)
    plain = np.zeros(m0.n)
    m0.apply(uc, plain)
    m1.linearize(np.ascontiguousarray(A["nw_us"]))
    mg.set_coarse_mass("galerkin")
    got = np.zeros(m0.n)
    m0.apply(uc, got)
    # by hand: M_1 w = J w - A w on the fine level
    w = np.zeros(m1.n)
    mg.prolong(0, uc, w)
    Jw = np.zeros(m1.n)
    m1.apply(w, Jw)
    m1.linearize(None)
    Aw = np.zeros(m1.n)
    m1.apply(w, Aw)
    t = np.zeros(m0.n)
    mg.restrict(0, Jw - Aw, t)
    assert rel(got - plain, t) < 1e-10
    # device-pointer call agrees
    d_uc = torch.from_numpy(uc).cuda()
    d_out = torch.empty_like(d_uc)
    m1.linearize(np.ascontiguousarray(A["nw_us"]))
    m0.apply(d_uc, d_out)
    m0.sync()
    assert np.array_equal(d_out.cpu().numpy(), got)
    m1.linearize(None)
    mg.set_coarse_mass(False)
    mg.close()
    m0.close()


def _hierarchy(A, m1):
    from disco4est_b200 import Mesh
    from disco4est_b200.runtime import Multigrid, cheby_params
    from disco4est_b200.sphere import sphere_desc

    ch = A["mg_cheby_params"]
    T0, g0 = sphere_desc(0, p=2, prefactor=2.0, radii=(1.0, 2.0, 3.0))
    m0 = Mesh(_robin(T0, 3.0), device=0, geom=g0)
    tr = [(8 * np.arange(13, dtype=np.int32), np.full(13, 8, np.int32))]
    sm = cheby_params(int(ch[0]), int(ch[1]), ch[2], ch[3], use_new=int(ch[4]))
    mg = Multigrid([m0, m1], tr, sm, vcycle_imax=1, vcycle_rtol=1e-9, bottom_solver="cheby",
                   bottom_cheby=cheby_params(15, 15, 30.0, 1.0))
    return m0, mg


@pytest.mark.parametrize("qtable", ["1", "0"])
def test_first_preconditioner_application_matches_reference(problem, qtable, monkeypatch):
    """y = B(x): input and output of the reference's first d4est_krylov_pc_multigrid_apply of the Newton solve
    (matrix operator on the coarse level as the reference EXECUTES it, mode "reference"); the Galerkin product
    gives another preconditioner"""
    # D4B_MG_QTABLE=0: the general (mixed-degree) form of the reference's slices instead of the per-level table
    monkeypatch.setenv("D4B_MG_QTABLE", qtable)
    A, m1 = problem
    m0, mg = _hierarchy(A, m1)
    m1.linearize(np.zeros(m1.n))
    mg.set_coarse_mass("reference")
    y = np.zeros(m1.n)
    mg.solve(y, np.ascontiguousarray(A["nw_pc_x0"]))
    err = rel(y, A["nw_pc_y0"])
    mg.set_coarse_mass("galerkin")
    y2 = np.zeros(m1.n)
    mg.solve(y2, np.ascontiguousarray(A["nw_pc_x0"]))
    m1.linearize(None)
    mg.close()
    m0.close()
    assert err < 1e-9
    assert rel(y2, A["nw_pc_y0"]) > 1e-3


def test_newton_iteration_counts_match_reference(problem):
    """d4b_newton_solve: same Newton steps, same FCG iterations in every step, same residual norms and solution as
    d4est_solver_newton_solve + d4est_solver_fcg_solve + d4est_krylov_pc_multigrid in the reference"""
    from disco4est_b200 import Mesh
    from disco4est_b200.runtime import Multigrid, cheby_params
    from disco4est_b200.sphere import sphere_desc

    A, m1 = problem
    prm = A["nw_params"]
    m0, mg = _hierarchy(A, m1)
    mg.set_coarse_mass("reference")
    u = np.zeros(m1.n)
    ierr, steps, its, norms, japplies = m1.newton_solve(u, pc=mg, atol=prm[0], rtol=prm[1], imin=int(prm[2]),
                                                        imax=int(prm[3]), fcg_atol=prm[4], fcg_rtol=prm[5],
                                                        fcg_imax=int(prm[6]))
    print("device Newton: steps %d, FCG iterations %s, |F| %s" % (steps, its, ["%.3e" % x for x in norms]))
    mg.close()
    m0.close()
    ref_its = A["nw_fcg_iterations"].astype(int).tolist()
    assert ierr == 0
    assert steps == len(ref_its)
    assert its == ref_its
    ref_norms = A["nw_residual_norms"]
    for a, b in zip(norms, ref_norms):
        assert abs(a - b) <= 1e-6 * b + 1e-11 * ref_norms[0]   # the last norms are round-off of the first
    assert rel(u, A["nw_u"]) < 1e-9
    assert japplies == sum(ref_its) + len(ref_its)
