#!/bin/bash
# final one-GPU run of the round: smoke, default bench (after the hp kernel changes)
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02_bench_n1_final2.json 2> gpurun_out/r02_bench_n1_final2.err
tail -c 400 gpurun_out/r02_bench_n1_final2.err
python - <<'P'
import json
for l in open('gpurun_out/r02_bench_n1_final2.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d.get('clocks'))
        for k,v in d.get('cubed_sphere',{}).items():
            if isinstance(v,dict) and 'GDOF_per_s' in v: print(k, round(v['GDOF_per_s'],2))
P
