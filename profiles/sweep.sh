#!/bin/bash
# degree sweep of the single-GPU apply; writes one JSON line per degree
out=${1:-gpurun_out/sweep.jsonl}
: > $out
for d in 1 2 3 4 5 6 7; do
  lvl=6; if [ $d -ge 4 ]; then lvl=5; fi
  python bench.py --level $lvl --deg $d --steps 20 --warmup 5 --no-cpu --no-vcycle --no-sphere 2>/dev/null | tail -1 >> $out
done
python - "$out" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    l = l.strip()
    if not l.startswith("{"): continue
    j = json.loads(l)
    print(j["config"]["workload"].split(",")[1:3], "GDOF/s %.2f" % j["value"], "ms %.4f" % j["ms_per_step"], "frac %.3f" % j["roofline"]["frac"], "e2e %.2f" % j["e2e"]["value"])
PY
