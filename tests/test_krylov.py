"""Outer Krylov loops: the reference's in-house CG and multigrid-preconditioned
flexible CG, device-resident.  tests/golden/brick_mg_l2_p2 holds the iterates and iteration counts of
the UNMODIFIED d4est_solver_fcg_solve / d4est_solver_cg_solve (oracle/ref_driver.c --fcg)."""
import numpy as np
import pytest

from oracle_lib import Oracle
from solver_ref import Transfer, cg, fcg, vcycle

CASE = "brick_mg_l2_p2"


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def hierarchy(A):
    from disco4est_b200.brick import brick_hierarchy

    return brick_hierarchy(2, int(A["deg"][0]))


def test_restatement_matches_reference_krylov(golden):
    """pins tests/solver_ref.py:cg / fcg (on the pinned oracle apply and V-cycle) to the reference"""
    A = golden(CASE)
    O = Oracle(A)
    rhs = np.ascontiguousarray(A["mg_rhs"])
    imax, rtol, atol, its_ref = A["cg_params"]
    u, its = cg(O.apply, np.zeros(rhs.size), rhs, int(imax), rtol, atol)
    assert its == int(its_ref)
    assert rel(u, A["cg_u"]) < 1e-9
    imax, rtol, atol, its_ref = A["fcg_params"]
    u, its, _ = fcg(O.apply, None, np.zeros(rhs.size), rhs, int(imax), rtol, atol)
    assert its == int(A["fcg_nopc_iters"][0])
    assert rel(u, A["fcg_nopc_u"]) < 1e-9
    # multigrid-preconditioned: one V-cycle from a zero guess per application
    descs, transfers = hierarchy(A)
    orcs = [Oracle(d) for d in descs]
    trs = [Transfer(descs[l], descs[l + 1], *transfers[l]) for l in range(len(descs) - 1)]
    cimax, eimax, ratio, mult, use_new = (float(x) for x in A["mg_cheby_params"])

    def pc(r):
        return vcycle([o.apply for o in orcs], trs, np.zeros(r.size), r, int(cimax), int(eimax), ratio, mult, int(use_new))[0]

    u, its, hist = fcg(O.apply, pc, np.zeros(rhs.size), rhs, int(imax), rtol, atol)
    assert its == int(its_ref) == len(A["fcg_r2"])
    assert np.allclose(hist, A["fcg_r2"], rtol=1e-5)
    assert rel(u, A["fcg_u"]) < 1e-9


@pytest.fixture()
def gpu_case(golden):
    from disco4est_b200 import Mesh
    from disco4est_b200.runtime import Multigrid, cheby_params

    A = golden(CASE)
    descs, transfers = hierarchy(A)
    meshes = [Mesh(d) for d in descs]
    cimax, eimax, ratio, mult, use_new = (float(x) for x in A["mg_cheby_params"])
    mg = Multigrid(meshes, transfers, cheby_params(int(cimax), int(eimax), ratio, mult, int(use_new)), vcycle_imax=1,
                   vcycle_rtol=1e-15)
    yield A, meshes[-1], mg
    mg.close()
    for m in meshes:
        m.close()


@pytest.mark.gpu
def test_gpu_cg_matches_reference(gpu_case):
    A, mesh, _ = gpu_case
    rhs = np.ascontiguousarray(A["mg_rhs"])
    imax, rtol, atol, its_ref = A["cg_params"]
    u = np.zeros(rhs.size)
    its = mesh.cg_solve(u, rhs, int(imax), rtol, atol)
    assert its == int(its_ref)
    assert rel(u, A["cg_u"]) < 1e-9
    # iteration cap is honoured exactly
    u2 = np.zeros(rhs.size)
    assert mesh.cg_solve(u2, rhs, 7, rtol, atol) == 7


@pytest.mark.gpu
def test_gpu_fcg_matches_reference(gpu_case):
    A, mesh, mg = gpu_case
    rhs = np.ascontiguousarray(A["mg_rhs"])
    imax, rtol, atol, its_ref = A["fcg_params"]
    # identity preconditioner (do_not_use_preconditioner = 1)
    u = np.zeros(rhs.size)
    its, _ = mesh.fcg_solve(u, rhs, int(imax), rtol, atol)
    assert its == int(A["fcg_nopc_iters"][0])
    # (without a preconditioner the executed recurrence does not contract -- it runs into imax with a
    #  growing iterate -- so the different summation order of the dots is amplified: 1e-9 measured)
    assert rel(u, A["fcg_nopc_u"]) < 1e-7
    # device V-cycle as the preconditioner: same iteration count, residual history and iterate
    u = np.zeros(rhs.size)
    its, hist = mesh.fcg_solve(u, rhs, int(imax), rtol, atol, pc=mg)
    assert its == int(its_ref)
    assert np.allclose(hist, A["fcg_r2"], rtol=1e-5)
    assert rel(u, A["fcg_u"]) < 1e-9
    # the recurrence of the reference's comments (D4B_FCG_TEXTBOOK) converges where the executed one stalls
    O = Oracle(A)
    u = np.zeros(rhs.size)
    its, _ = mesh.fcg_solve(u, rhs, 200, rtol, atol, textbook=True)
    u_want, its_want, _ = fcg(O.apply, None, np.zeros(rhs.size), rhs, 200, rtol, atol, textbook=True)
    assert abs(its - its_want) <= 1 and rel(u, u_want) < 1e-8
    u = np.zeros(rhs.size)
    # device pointers: same bits
    import torch

    its, _ = mesh.fcg_solve(u, rhs, int(imax), rtol, atol, pc=mg)
    du = torch.zeros(rhs.size, dtype=torch.float64, device="cuda")
    its2, _ = mesh.fcg_solve(du, torch.from_numpy(rhs).cuda(), int(imax), rtol, atol, pc=mg)
    assert its2 == its and np.array_equal(du.cpu().numpy(), u)


@pytest.mark.gpu
def test_blas1_kernels_direct(golden):
    """d4b_vec_axpy / xpby / scale / dot against numpy (d4est_linalg_vec_axpy: y += a x; _xpby: y = x + b y;
    _scale: x *= a; src/LinearAlgebra/d4est_linalg.c:179-236), a length that is not a multiple of any block size"""
    import ctypes as C

    import torch
    from disco4est_b200 import Mesh

    m = Mesh(golden(CASE))
    n = 100003
    rng = np.random.default_rng(3)
    x, y = rng.standard_normal(n), rng.standard_normal(n)
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    L, h = m.L, m.h
    L.d4b_vec_axpy(h, 0.37, C.c_void_p(dx.data_ptr()), C.c_void_p(dy.data_ptr()), n)
    m.sync()
    y = y + 0.37 * x
    assert np.allclose(dy.cpu().numpy(), y, rtol=1e-14, atol=1e-14)   # (FMA on the device, separate multiply and add in numpy)
    L.d4b_vec_xpby(h, C.c_void_p(dx.data_ptr()), -1.25, C.c_void_p(dy.data_ptr()), n)
    m.sync()
    y = x - 1.25 * y
    assert np.allclose(dy.cpu().numpy(), y, rtol=1e-14, atol=1e-14)   # (FMA on the device, separate multiply and add in numpy)
    L.d4b_vec_scale(h, 3.0, C.c_void_p(dy.data_ptr()), n)
    m.sync()
    y = 3.0 * y
    assert np.allclose(dy.cpu().numpy(), y, rtol=1e-14, atol=1e-14)   # (FMA on the device, separate multiply and add in numpy)
    d = float(L.d4b_vec_dot(h, C.c_void_p(dx.data_ptr()), C.c_void_p(dy.data_ptr()), n))
    assert abs(d - float(x @ y)) <= 1e-12 * float(np.abs(x) @ np.abs(y))
    m.close()
