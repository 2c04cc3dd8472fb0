#!/usr/bin/env python
"""Two-puncture hierarchy (13-tree sphere, Robin boundary, level argv[1]): seconds per V-cycle without a mass term,
with the linearised term on the finest level only, and with the coarse-level matrix operator (Galerkin / reference)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from disco4est_b200 import Mesh
from disco4est_b200.runtime import Multigrid, cheby_params
from disco4est_b200.sphere import sphere_desc
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import PUNCTURES

nl = int(sys.argv[1]) if len(sys.argv) > 1 else 4
R2 = 3.0
def robin(A):
    mp = int(np.asarray(A["local_mortar_nodes_quad"]).reshape(-1)[0])
    A = dict(A); A["bc_type"] = 1
    A["robin_coeff_mortar_quad"] = np.full(mp, 1.0 / R2); A["robin_rhs_mortar_quad"] = np.zeros(mp)
    return A
meshes = []
for l in range(nl + 1):
    A, g = sphere_desc(l, p=3, prefactor=2.0, radii=(1.0, 2.0, R2))
    meshes.append(Mesh(robin(A), geom=g))
transfers = [(8 * np.arange(13 * 8 ** l, dtype=np.int32), np.full(13 * 8 ** l, 8, np.int32)) for l in range(nl)]
top = meshes[-1]
mg = Multigrid(meshes, transfers, cheby_params(8, 10, 30.0, 1.0), vcycle_imax=1, vcycle_rtol=1e-9,
               bottom_solver="cheby", bottom_cheby=cheby_params(15, 15, 30.0, 1.0))
rhs = torch.rand(top.n, dtype=torch.float64, device="cuda")
u = torch.zeros_like(rhs)
def timed(tag):
    for _ in range(4): mg.vcycle(u, rhs)
    torch.cuda.synchronize(); t0 = time.perf_counter(); r0 = mg.graph_replays()
    for _ in range(5): mg.vcycle(u, rhs)
    torch.cuda.synchronize()
    print("%-46s %.2f ms per V-cycle (%d of 5 replayed)" % (tag, (time.perf_counter() - t0) / 5 * 1e3, mg.graph_replays() - r0), flush=True)
timed("no mass term")
top.set_nonlinear_term("punctures", punctures=PUNCTURES)
top.linearize(0.05 * torch.rand(top.n, dtype=torch.float64, device="cuda"))
timed("linearised term on the finest level only")
for mode in ("galerkin", "reference"):
    mg.set_coarse_mass(mode)
    timed("coarse-level matrix operator: " + mode)
v = torch.rand(top.n, dtype=torch.float64, device="cuda"); w = torch.empty_like(v)
for _ in range(3): top.apply(v, w)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): top.apply(v, w)
torch.cuda.synchronize(); print("Jacobian apply on the finest level: %.3f ms" % ((time.perf_counter() - t0) / 20 * 1e3))
