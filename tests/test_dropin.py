"""Drop-in check: the UNMODIFIED reference (oracle/_ref/ref_driver_b200 = reference sources +
adapter + libd4est_b200.so) calling eqns->apply_lhs = d4est_b200_apply_lhs -- directly and
from inside its own multigrid solver.  Skipped when the binary was not built (it is built
by __graft_entry__.build() in the container that has /root/reference)."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRV = os.path.join(ROOT, "oracle", "_ref", "ref_driver_b200")
CFG = os.path.join(ROOT, "oracle", "configs")


def run(ini, *args):
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
    out = subprocess.run([DRV, os.path.join(CFG, ini), "-", *args, "--b200"], env=env, capture_output=True,
                         text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    return [json.loads(l) for l in out.stdout.splitlines() if l.startswith("{")]


pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(DRV), reason="ref_driver_b200 not built")]


@pytest.mark.parametrize("ini,args", [("brick_poisson.ini", ["--uniform-deg", "3"]),
                                      ("sphere_p.ini", ["--uniform-deg", "2"]),
                                      ("sphere_hp.ini", []),
                                      ("sphere_hp.ini", ["--amr", "1"]),
                                      ("brick_poisson.ini", ["--uniform-deg", "2", "--bc", "robin"]),
                                      ("sphere_p.ini", ["--amr", "1", "--bc", "robin"])])
def test_apply_lhs_callback_inside_the_reference(ini, args):
    lines = run(ini, *args)
    res = [l for l in lines if "b200_apply_rel_l2" in l][0]
    assert res["b200_apply_rel_l2"] < 1e-12


@pytest.mark.parametrize("ini,args", [("brick_poisson.ini", ["--uniform-deg", "3"]),
                                      ("sphere_p.ini", ["--uniform-deg", "2"]),
                                      ("sphere_p.ini", ["--amr", "1"]),
                                      ("sphere_hp.ini", ["--amr", "1"])])
def test_device_geometry_inside_the_reference(ini, args):
    """d4est_b200_ctx_use_device_geometry: the adapter hands the forest positions of the reference's
    elements to d4b_mesh_create_with_geometry; the operator equals d4est_laplacian_apply_aij"""
    lines = run(ini, *args)
    res = [l for l in lines if "b200_device_geometry_apply_rel_l2" in l][0]
    assert res["b200_device_geometry_apply_rel_l2"] < 1e-12


@pytest.mark.parametrize("ini,args", [("brick_poisson.ini", ["--uniform-deg", "2", "--bc", "poly"]),
                                      ("sphere_p.ini", ["--amr", "1", "--bc", "poly"]),
                                      ("sphere_hp.ini", ["--amr", "1", "--bc", "poly"])])
def test_build_rhs_inside_the_reference(ini, args):
    """d4est_b200_build_rhs_with_strong_bc against d4est_laplacian_build_rhs_with_strong_bc (same arguments),
    f on the LGL nodes and on the quadrature points, Dirichlet data evaluated by the adapter"""
    lines = run(ini, *args)
    res = [l for l in lines if "b200_build_rhs_rel_l2" in l][0]
    assert res["b200_build_rhs_rel_l2"] < 1e-12
    # d4est_b200_estimator_bi_compute against d4est_estimator_bi_compute, same arguments
    res = [l for l in lines if "b200_estimator_rel_max" in l][0]
    assert res["b200_estimator_rel_max"] < 1e-11


@pytest.mark.parametrize("ini,args", [("brick_poisson.ini", ["--uniform-deg", "3"]), ("sphere_hp.ini", ["--amr", "1"])])
def test_jacobian_apply_with_mass_term_inside_the_reference(ini, args):
    """d4est_b200_set_mass_term + apply_lhs against apply_aij + the reference's fofufofvlilj loop"""
    lines = run(ini, *args)
    res = [l for l in lines if "b200_jacobian_apply_rel_l2" in l][0]
    assert res["b200_jacobian_apply_rel_l2"] < 1e-12


def test_reference_multigrid_driving_our_apply():
    lines = run("brick_poisson_mg.ini", "--uniform-deg", "3", "--mg", "4")
    res = [l for l in lines if "r2_b200_apply" in l][0]
    assert res["u_rel_l2_diff"] < 1e-12
    assert res["r2_worst_rel_diff"] < 1e-6
    assert res["device_mesh_uploads"] == 3          # one upload per multigrid level, cached afterwards
    for a, b in zip(res["r2_reference"], res["r2_b200_apply"]):
        assert abs(a - b) <= 1e-6 * a


def test_reference_multigrid_with_our_smoother_entry():
    """d4est_b200_smoother_cheby_init installed as mg_data->smoother of the reference's solver"""
    lines = run("brick_poisson_mg.ini", "--uniform-deg", "3", "--mg", "4")
    res = [l for l in lines if "smoother_u_rel_l2_diff" in l][0]
    assert res["smoother_u_rel_l2_diff"] < 1e-10
    assert res["smoother_r2_worst_rel_diff"] < 1e-5


@pytest.mark.parametrize("ini,args", [("brick_poisson_mg.ini", ["--uniform-deg", "3", "--mg", "4"]),
                                      ("sphere_hp_mg.ini", ["--amr", "1", "--mg", "3"]),
                                      ("sphere_p_mg.ini", ["--amr", "1", "--mg", "3"])])
def test_reference_preconditioner_slot(ini, args):
    """d4est_b200_pc_multigrid_create against d4est_krylov_pc_multigrid_create, both called as
    pc->pc_apply(pc, x, y): hierarchy recorded from the reference's own V-cycle, V-cycles on the device
    (brick: fused smoother, bottom CG; hp sphere: general kernel, hp transfers, bottom Chebyshev; hp sphere with
    deg_quad = deg: mortar kernel / face split with the fused smoother on every level)"""
    lines = run(ini, *args)
    res = [l for l in lines if "pc_y_rel_l2_diff" in l][0]
    assert res["pc_vcycles_b200"] == res["pc_vcycles_reference"]
    assert res["pc_y_rel_l2_diff"] < 1e-9
    assert res["pc_hierarchy_builds"] == 1


def test_reference_fcg_with_our_preconditioner():
    """the UNMODIFIED d4est_solver_fcg_solve with pc = d4est_b200_pc_multigrid_create and
    apply_lhs = d4est_b200_apply_lhs: same Krylov iteration count and solution as the stock run"""
    lines = run("brick_poisson_mg.ini", "--uniform-deg", "2", "--mg", "1", "--fcg", "30")
    res = [l for l in lines if "fcg_iters_b200_pc" in l][0]
    assert res["fcg_iters_b200_pc"] == res["fcg_iters_reference"]
    assert res["fcg_u_rel_l2_diff"] < 1e-9


@pytest.mark.parametrize("ini", ["sphere_p.ini", "sphere_hp.ini"])
def test_cached_mesh_is_dropped_after_count_preserving_p_adapt(ini):
    """The reference reuses one d4est_mesh_data_t across d4est_mesh_update; two elements exchanging their
    degrees keep every count.  The adapter's cache key (stride hash + forest hash) must notice it."""
    lines = run(ini, "--amr", "1")
    res = [l for l in lines if "b200_apply_after_degree_swap_rel_l2" in l][0]
    assert res["same_counts"] == 1
    assert res["reference_change_rel_l2"] > 1e-6      # the operator really changed
    assert res["uploads_for_swapped_mesh"] == 1
    assert res["b200_apply_after_degree_swap_rel_l2"] < 1e-12


def test_reference_newton_with_device_residual_jacobian_and_preconditioner():
    """BASELINE configs[4] in small, INSIDE the reference: its unmodified d4est_solver_newton_solve and
    d4est_solver_fcg_solve run with build_residual = d4est_b200_build_residual, apply_lhs = d4est_b200_apply_lhs
    (Jacobian about prob_vecs->u0) and pc = d4est_b200_pc_multigrid_create; puncture terms are device functors.
    Same Newton steps and FCG iterations per step as the stock run, one upload of u0 per Newton step."""
    lines = run("sphere_punctures.ini", "--newton")
    ref = [l for l in lines if "newton_steps" in l][0]
    res = [l for l in lines if "b200_newton_steps" in l][0]
    assert res["b200_newton_ierr"] == 0
    assert res["b200_newton_steps"] == ref["newton_steps"]
    assert res["b200_fcg_iterations_per_step"] == ref["fcg_iterations_per_step"]
    assert res["same_iteration_counts"] == 1
    assert res["residual_norm_worst_rel_diff"] < 1e-4   # the last norms (1e-9 of the first) are round-off of the first
    assert res["newton_u_rel_l2_diff"] < 1e-9
    assert res["u0_uploads"] == res["b200_newton_steps"]
