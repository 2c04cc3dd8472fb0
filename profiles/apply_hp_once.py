#!/usr/bin/env python
"""Applies on the cubed sphere with p-jumps (hp mortar kernel): seeded random degree 1..pmax per element and the
radially graded mesh; device-timed over 20 applies, with a checksum so that two library builds can be compared
(D4B_LIBRARY_PATH).  usage: apply_hp_once.py [level] [pmax]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from disco4est_b200 import Mesh
from disco4est_b200.sphere import sphere_desc

level = int(sys.argv[1]) if len(sys.argv) > 1 else 5
pmax = int(sys.argv[2]) if len(sys.argv) > 2 else 3
tag = os.path.basename(os.environ.get("D4B_LIBRARY_PATH", "base"))
for name, kw in (("random_hp", dict(pmax=pmax)), ("graded_hp", dict(p=pmax, graded=True))):
    A, g = sphere_desc(level, **kw)
    m = Mesh(A, geom=g)
    st = torch.cuda.current_stream()
    m.set_stream(st.cuda_stream)
    u = torch.rand(m.n, dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(7))
    Au = torch.zeros_like(u)
    for _ in range(3):
        m.apply(u, Au)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(20):
        m.apply(u, Au)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print("%-16s %s level %d pmax %d  hp elements %d  %.4f ms  %.2f GDOF/s  checksum %.10e" %
          (tag, name, level, pmax, m.hp_elements(), ms, m.n / ms / 1e6, float(Au.double().sum())), flush=True)
    m.close()
