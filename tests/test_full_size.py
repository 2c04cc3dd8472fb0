"""BASELINE.json configs[1] at full size (64^3-element unit cube, p=3, 16.8 M DOF) through size-independent
properties: the reference's own known answer A(x^2+y^2+z^2) = M(-6)
(src/Tests/Unit/d4est_test_laplacian_consistency.c:21,418-426), symmetry, linearity, and the pipelined
host-pointer apply against the device-pointer apply."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cube():
    import torch

    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_desc, brick_xyz

    level, p = 6, 3
    m = Mesh(brick_desc(level, p))
    xyz = brick_xyz(level, p)
    yield m, xyz, level, p, torch
    m.close()


def test_polynomial_known_answer(cube):
    from disco4est_b200.brick import lgl_nodes

    m, xyz, level, p, torch = cube
    N, n = p + 1, 1 << level
    h = 1.0 / n
    u = np.ascontiguousarray((xyz ** 2).sum(axis=0))
    Au = np.zeros_like(u)
    m.apply(u, Au)
    # int phi_i = LGL weight (the N-point Gauss rule of the operator integrates phi_i exactly)
    x = lgl_nodes(N)
    P = np.polynomial.legendre.Legendre.basis(p)
    w = 2.0 / (p * (p + 1) * P(x) ** 2)
    w3 = (w[None, None, :] * w[None, :, None] * w[:, None, None]).reshape(-1) * (0.5 * h) ** 3
    A3 = Au.reshape(-1, N ** 3)
    # elements that do not touch the boundary: u is a continuous polynomial of degree 2 <= p, all jumps
    # vanish and the homogeneous Dirichlet data is never seen
    lo, hi = xyz.reshape(3, -1, N ** 3).min(axis=2), xyz.reshape(3, -1, N ** 3).max(axis=2)
    interior = ((lo > 1e-12) & (hi < 1 - 1e-12)).all(axis=0)
    assert interior.sum() == (n - 2) ** 3
    err = np.abs(A3[interior] + 6.0 * w3[None, :]).max()
    assert err < 1e-10 * (6.0 * w3.max())


def test_symmetry_and_linearity(cube):
    m, xyz, level, p, torch = cube
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(9)
    u = torch.randn(m.n, dtype=torch.float64, device=dev, generator=g)
    v = torch.randn(m.n, dtype=torch.float64, device=dev, generator=g)
    Au, Av, Auv = torch.empty_like(u), torch.empty_like(u), torch.empty_like(u)
    m.apply(u, Au); m.apply(v, Av); m.apply(u + 2.0 * v, Auv)
    m.sync()
    a, b = float(v.dot(Au)), float(u.dot(Av))
    assert abs(a - b) < 1e-11 * abs(a)
    assert float((Auv - Au - 2.0 * Av).norm()) < 1e-12 * float(Auv.norm())
    assert float(u.dot(Au)) > 0.0


def test_pipelined_host_apply_full_size(cube):
    m, xyz, level, p, torch = cube
    u = torch.rand(m.n, dtype=torch.float64).pin_memory()
    Au = torch.empty(m.n, dtype=torch.float64).pin_memory()
    m.apply(u.numpy(), Au.numpy())
    du = u.cuda()
    dAu = torch.empty_like(du)
    m.apply(du, dAu)
    m.sync()
    assert torch.equal(Au, dAu.cpu())
