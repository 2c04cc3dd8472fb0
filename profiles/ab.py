#!/usr/bin/env python
"""A/B timing of environment-variable variants of the apply, interleaved.

usage: python profiles/ab.py [--level 6 --deg 3 --rounds 4] VAR=a,b[,c] ...
Each variant runs in a fresh process (the env is read at mesh creation);
prints the per-variant min/median ms over the rounds.
"""
import itertools, json, os, subprocess, sys
args = sys.argv[1:]
level, deg, rounds = 6, 3, 4
sweep = []
i = 0
while i < len(args):
    if args[i] == "--level": level = int(args[i + 1]); i += 2
    elif args[i] == "--deg": deg = int(args[i + 1]); i += 2
    elif args[i] == "--rounds": rounds = int(args[i + 1]); i += 2
    else:
        k, v = args[i].split("="); sweep.append((k, v.split(","))); i += 1
variants = list(itertools.product(*[[(k, x) for x in vs] for k, vs in sweep])) or [()]
res = {v: [] for v in variants}
for r in range(rounds):
    for v in variants:
        env = dict(os.environ); env.update({k: x for k, x in v})
        out = subprocess.run([sys.executable, "bench.py", "--level", str(level), "--deg", str(deg), "--steps", "30",
                              "--warmup", "5", "--no-cpu", "--no-vcycle", "--no-sphere"], env=env, capture_output=True, text=True).stdout
        j = json.loads(out.strip().splitlines()[-1])
        res[v].append(j["ms_per_step"])
for v, ms in res.items():
    ms = sorted(ms)
    print(dict(v), "min %.4f med %.4f ms  -> %.2f GDOF/s (best)" % (ms[0], ms[len(ms) // 2], j["config"]["dof_per_gpu"] / ms[0] / 1e6))
