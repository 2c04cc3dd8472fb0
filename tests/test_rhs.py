"""Right-hand side with strongly imposed boundary data (first half):
d4est_laplacian_build_rhs_with_strong_bc (src/dGMath/d4est_laplacian.c:15-140), rhs = M f - A(0).
The fixtures with inhomogeneous Dirichlet data hold f on the LGL nodes and on the quadrature points
and the rhs the unmodified reference built from each."""
import numpy as np
import pytest

CASES = ["brick_l2_p2_bc", "sphere_hp_amr0", "sphere_hp_amr1", "sphere_p_amr1"]
TOL = 1e-12


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("case", CASES)
def test_oracle_composition_matches_reference(golden, case):
    """the oracle's operators composed the way the reference composes its own"""
    import oracle_lib

    A = golden(case)
    O = oracle_lib.Oracle(A)
    J = np.ascontiguousarray(A["J_quad"])
    A0 = O.apply(np.zeros(O.n))
    assert np.linalg.norm(A0) > 0          # the Dirichlet data is not zero
    assert rel(O.apply_mass(A["rhs_f_lobatto"], J) - A0, A["rhs_strong_bc_lobatto"]) < TOL
    assert rel(O.apply_galerkin(A["rhs_f_quad"], J) - A0, A["rhs_strong_bc_quad"]) < TOL


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_device_build_rhs_matches_reference(golden, case):
    from disco4est_b200 import Mesh

    A = golden(case)
    m = Mesh(A)
    J = np.ascontiguousarray(A["J_quad"])
    rhs = np.zeros(m.n)
    m.build_rhs(np.ascontiguousarray(A["rhs_f_lobatto"]), J, rhs)
    assert rel(rhs, A["rhs_strong_bc_lobatto"]) < TOL
    m.build_rhs(np.ascontiguousarray(A["rhs_f_quad"]), J, rhs, on_quad=True)
    assert rel(rhs, A["rhs_strong_bc_quad"]) < TOL
    m.close()


@pytest.mark.gpu
def test_device_build_rhs_device_pointers(golden):
    import torch

    from disco4est_b200 import Mesh

    A = golden("sphere_p_amr1")
    m = Mesh(A)
    dev = torch.device("cuda:0")
    f = torch.from_numpy(np.ascontiguousarray(A["rhs_f_lobatto"])).to(dev)
    J = torch.from_numpy(np.ascontiguousarray(A["J_quad"])).to(dev)
    rhs = torch.zeros(m.n, dtype=torch.float64, device=dev)
    m.build_rhs(f, J, rhs)
    assert rel(rhs.cpu().numpy(), A["rhs_strong_bc_lobatto"]) < TOL
    m.close()


# ---------------------------------------------------------------- estimator (second half of f4)
@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("family,key", [("houston_maxp_minh", "houston"), ("bi_max_p2_over_h", "bi_p2h")])
def test_device_estimator_matches_reference(golden, case, family, key):
    """d4b_estimator_bi against d4est_estimator_bi_compute of the unmodified reference: eta^2 of every
    element and its four parts (residual, Je1, Je2, boundary Je2), two prefactor families"""
    from disco4est_b200 import Mesh

    A = golden(case)
    m = Mesh(A)
    c = lambda k: np.ascontiguousarray(A[k])
    nl = int(A["n_local"][0])
    terms = np.zeros(4 * nl)
    eta2 = m.estimator_bi(c("u"), c("est_residual"), c("J_quad"), c("est_diam_volume"), c("sj_m_mortar_quad"),
                          c("hm_mortar_quad"), c("hp_mortar_quad"), float(A["est_penalty_prefactor"][0]), family=family,
                          dirichlet_g=c("dirichlet_g_mortar_quad"), terms=terms)
    m.close()
    ref, ref_terms = A["est_eta2_" + key], A["est_terms_" + key].reshape(4, nl)
    assert np.abs(eta2 - ref).max() < 1e-11 * np.abs(ref).max()
    t = terms.reshape(4, nl)
    for i in range(4):
        assert np.abs(t[i] - ref_terms[i]).max() < 1e-11 * max(np.abs(ref_terms[i]).max(), 1e-300), i
