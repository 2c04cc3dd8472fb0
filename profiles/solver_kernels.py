#!/usr/bin/env python
"""One cg_eigs(5) and one Chebyshev smoothing (5 sweeps) on a 2^level cube, p=3: for an ncu launch list
(per-kernel durations of the solver sequences)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from disco4est_b200 import Mesh
from disco4est_b200.brick import brick_desc

level = int(sys.argv[1]) if len(sys.argv) > 1 else 6
A, g = brick_desc(level, 3, geometry=False)
m = Mesh(A, geom=g)
rhs = torch.rand(m.n, dtype=torch.float64, device="cuda")
u = torch.zeros_like(rhs); r = torch.zeros_like(rhs)
lmax = m.cg_eigs(u, rhs, 5)
m.cheby_smooth(u, rhs, r, 5, lmax / 30, lmax)
m.sync()
print("lmax", lmax, float(r.norm()))
