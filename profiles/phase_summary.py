#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` dump into barrier-delimited phases.

usage: ncu -i X.ncu-rep --page source --csv --kernel-name regex:K > src.csv
       python profiles/phase_summary.py src.csv
Prints, per phase (SASS between two BAR.SYNC), warp instructions, stall samples by
reason, shared-memory wavefronts (actual / ideal) and the opcode mix.
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r][0]
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_")]


def num(r, name):
    try:
        return float(r[col[name]] or 0)
    except ValueError:
        return 0.0


phases = [collections.defaultdict(float)]
ops = [collections.Counter()]
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    sass = r[col["Source"]]
    ph, oc = phases[-1], ops[-1]
    inst = num(r, "Instructions Executed")
    ph["inst"] += inst
    ph["samples"] += num(r, "# Samples")
    ph["wf"] += num(r, "L1 Wavefronts Shared")
    ph["wf_ideal"] += num(r, "L1 Wavefronts Shared Ideal")
    ph["l2sec"] += num(r, "L2 Theoretical Sectors Global")
    for s in stalls:
        ph[s] += num(r, s)
    toks = sass.split()
    op = next((t for t in toks if not t.startswith("@")), "?").split(".")[0]
    oc[op] += inst
    if "BAR.SYNC" in sass:
        phases.append(collections.defaultdict(float))
        ops.append(collections.Counter())

tot = sum(p["samples"] for p in phases) or 1.0
for i, (p, o) in enumerate(zip(phases, ops)):
    top = sorted(((p[s], s) for s in stalls), reverse=True)[:4]
    print("phase %2d inst %9.2fM samples %5.1f%% smem wf %7.2fM/%7.2fM l2sec %7.2fM | %s" % (
        i, p["inst"] / 1e6, 100 * p["samples"] / tot, p["wf"] / 1e6, p["wf_ideal"] / 1e6, p["l2sec"] / 1e6,
        " ".join("%s %.0f%%" % (s[6:], 100 * v / max(p["samples"], 1)) for v, s in top)))
    print("          " + " ".join("%s %.1fM" % (k, v / 1e6) for k, v in o.most_common(12)))
