#!/usr/bin/env python
"""A few applies on the 13-tree cubed sphere (uniform degree, curved conforming kernel): target of ncu, and a
device-timed figure (30 applies).  usage: apply_sphere_once.py [level] [p]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from disco4est_b200 import Mesh
from disco4est_b200.sphere import sphere_desc

level = int(sys.argv[1]) if len(sys.argv) > 1 else 5
p = int(sys.argv[2]) if len(sys.argv) > 2 else 3
A, g = sphere_desc(level, p=p)
m = Mesh(A, geom=g)
st = torch.cuda.current_stream()
m.set_stream(st.cuda_stream)
u = torch.rand(m.n, dtype=torch.float64, device="cuda")
Au = torch.zeros_like(u)
for _ in range(4):
    m.apply(u, Au)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(30):
    m.apply(u, Au)
e1.record(st)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 30
print("%s sphere level %d p %d  %.4f ms  %.2f GDOF/s" % (os.path.basename(os.environ.get("D4B_LIBRARY_PATH", "base")), level, p, ms, m.n / ms / 1e6), flush=True)
