import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = [
    "brick_l2_p1", "brick_l2_p2_bc", "brick_l2_p3", "brick_l2_p5", "brick_l2_p7",
    "sphere_p2", "sphere_hp_amr0", "sphere_hp_amr1", "sphere_hp_amr1_zero", "sphere_p_amr1",
    "brick_l2_p2_robin", "sphere_p_amr1_robin",
    # the other three penalty functions (+ region multipliers != 1), d4est_laplacian_flux_sipg.c:945-1005, :1165-1199
    "sphere_p_pen_maxpp1", "sphere_p_pen_meanp_mult", "sphere_p_pen_meanph",
]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import d4bio

    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = d4bio.load(os.path.join(GOLDEN, name + ".d4b.gz"))
        return cache[name]

    return get
