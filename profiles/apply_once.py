#!/usr/bin/env python
"""A few applies on a 2^level cube (per-point metric), p from argv: target of `ncu --set full`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("D4B_AFFINE", "0")
import torch
from disco4est_b200 import Mesh
from disco4est_b200.brick import brick_desc

level = int(sys.argv[1]) if len(sys.argv) > 1 else 6
p = int(sys.argv[2]) if len(sys.argv) > 2 else 3
A, g = brick_desc(level, p, geometry=False)
m = Mesh(A, geom=g)
u = torch.rand(m.n, dtype=torch.float64, device="cuda")
Au = torch.zeros_like(u)
for _ in range(4):
    m.apply(u, Au)
m.sync()
print("checksum", float(Au.sum()))
