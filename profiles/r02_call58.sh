#!/bin/bash
# 2 GPUs: multi-rank tests (hp meshes with ghost neighbours included) and the sphere block of the bench after the hp kernel changes
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_rank.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r02_multi_rank_hp.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-vcycle --no-affine --no-punctures --no-halo-ab --no-parity > gpurun_out/r02_bench_n2_hp.json 2> gpurun_out/r02_bench_n2_hp.err
tail -c 600 gpurun_out/r02_bench_n2_hp.err; python - <<'P'
import json
for l in open('gpurun_out/r02_bench_n2_hp.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'])
        for k,v in d.get('cubed_sphere',{}).items():
            if isinstance(v,dict) and 'GDOF_per_s' in v: print(k, round(v['GDOF_per_s'],2), v['checksum_Au'])
P
