"""Time one apply on the synthetic cubed sphere (uniform-p curved kernel, random-hp kernel).
usage: python profiles/hp_bench.py [level] [deg] [mode ...]   mode: uniform | hp | graded | generic (uniform mesh forced to the general kernel) | amr (brick, one octant refined)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disco4est_b200 import Mesh
from disco4est_b200.sphere import sphere_desc

level = int(sys.argv[1]) if len(sys.argv) > 1 else 4
deg = int(sys.argv[2]) if len(sys.argv) > 2 else 3
modes = sys.argv[3:] or ["uniform", "hp"]  # also: graded
dev = torch.device("cuda:0")
for mode in modes:
    kw = dict(pmax=deg) if mode == "hp" else dict(p=deg, graded=True) if mode == "graded" else dict(p=deg)
    if mode == "generic":
        os.environ["D4B_FORCE_GENERIC"] = "1"
    if mode == "amr":
        from disco4est_b200.brick import brick_amr_desc
        A, g = brick_amr_desc(level, deg)
    else:
        A, g = sphere_desc(level, **kw)
    m = Mesh(A, geom=g)
    os.environ.pop("D4B_FORCE_GENERIC", None)
    stream = torch.cuda.current_stream()
    m.set_stream(stream.cuda_stream)
    u = torch.rand(m.n, dtype=torch.float64, device=dev)
    Au = torch.empty_like(u)
    for _ in range(3):
        m.apply(u, Au)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 10
    e0.record(stream)
    for _ in range(steps):
        m.apply(u, Au)
    e1.record(stream)
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3 / steps
    print("%-8s level %d deg %d: %8d elements %9d DOF  fast %d flat %d hp %d  %.3f ms  %.3f GDOF/s  checksum %.10e"
          % (mode, level, deg, A["n_local"], m.n, m.fast_elements(), m.flat_elements(), m.hp_elements(), 1e3 * t, m.n / t / 1e9,
             float(Au.sum())), flush=True)
    m.close()
