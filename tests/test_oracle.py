"""Pins the CPU oracle (oracle/d4b_oracle*.c) against the unmodified reference.

The fixtures under tests/golden/ were written by the reference itself
(oracle/_ref/ref_driver, see oracle/make_golden.sh): inputs, Au of
d4est_laplacian_apply_aij and of d4est_laplacian_with_opt_apply_aij, and the
reference's 1-D operator tables.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib
from conftest import GOLDEN_CASES

TOL = 1e-12  # relative L2, BASELINE.json north_star


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_oracle_matches_reference_apply(golden, case):
    A = golden(case)
    O = oracle_lib.Oracle(A)
    Au = O.apply(A["u"])
    assert rel(Au, A["Au_plain"]) < TOL
    assert rel(Au, A["Au_with_opt"]) < TOL
    # the reference's own gate between its two variants is 1e-13 absolute
    # (d4est_test_laplacian_speedup.c:485); the oracle stays inside it as well
    assert np.abs(Au - A["Au_plain"]).max() < 1e-12


def test_reference_variants_agree(golden):
    for case in GOLDEN_CASES:
        A = golden(case)
        assert np.abs(A["Au_plain"] - A["Au_with_opt"]).max() < 1e-13


def test_oracle_stiffness_only(golden):
    A = golden("brick_l2_p3")
    O = oracle_lib.Oracle(A)
    assert rel(O.apply_stiffness(A["u"]), A["Ku"]) < TOL


def _tab(L, fn, *args, n):
    out = np.zeros(n)
    fn(*args, out.ctypes.data_as(C.POINTER(C.c_double)))
    return out


def test_tables_match_reference(golden):
    A = golden("brick_l2_p7")
    L = oracle_lib.lib()
    P = C.POINTER(C.c_double)
    for p in range(1, 8):
        N = p + 1
        x, w = np.zeros(N), np.zeros(N)
        L.d4bo_lobatto(N, x.ctypes.data_as(P), w.ctypes.data_as(P))
        assert np.abs(x - A["tab_lgl_nodes_%d" % p]).max() < 1e-14
        assert np.abs(w - A["tab_lgl_weights_%d" % p]).max() < 1e-14
        L.d4bo_gauss(N, x.ctypes.data_as(P), w.ctypes.data_as(P))
        assert np.abs(x - A["tab_gauss_nodes_%d" % p]).max() < 1e-14
        assert np.abs(w - A["tab_gauss_weights_%d" % p]).max() < 1e-14
        D = _tab(L, L.d4bo_dij, p, n=N * N)
        assert np.abs(D - A["tab_dij_%d" % p]).max() < 5e-13
        for q in range(p, 8):
            E = _tab(L, L.d4bo_eval_matrix, p, q, 0, n=(q + 1) * N)
            assert np.abs(E - A["tab_lgl_to_gauss_%d_%d" % (p, q)]).max() < 1e-13


def test_operator_symmetric(golden):
    """dense A = A^T (d4est_test_laplacian_symmetry.c:21, homogeneous BC)."""
    A = golden("sphere_hp_amr0")
    A = dict(A)
    A.pop("dirichlet_g_mortar_quad", None)
    O = oracle_lib.Oracle(A)
    n = O.n
    rng = np.random.default_rng(0)
    x, y = rng.standard_normal(n), rng.standard_normal(n)
    lhs, rhs = y @ O.apply(x), x @ O.apply(y)
    assert abs(lhs - rhs) / abs(lhs) < 1e-11
