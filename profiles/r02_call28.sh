cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_multi_rank.py -m gpu -q -x -v 2>&1 | tail -14) > gpurun_out/r02_multi_rank_tests_2gpu_final.txt
cat gpurun_out/r02_multi_rank_tests_2gpu_final.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_n2_final.json 2> gpurun_out/r02_bench_n2_final.err
echo "rc=$?"; tail -2 gpurun_out/r02_bench_n2_final.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n2_final.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
print('parity', d.get('parity',{}).get('nccl_rel_l2'))
print('ab', d.get('halo_comparison'))
print('vcycle', d.get('vcycle',{}).get('seconds_per_vcycle'))
a=d.get('affine_fast_path',{}); print('affine', a.get('GDOF_per_s'), a.get('vcycle',{}).get('seconds_per_vcycle'))
print('sphere', {k:round(v['GDOF_per_s'],2) for k,v in d['cubed_sphere'].items() if isinstance(v,dict)})
PY
