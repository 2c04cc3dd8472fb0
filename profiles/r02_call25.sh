cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02_bench_n8_final.json 2> gpurun_out/r02_bench_n8_final.err
echo "rc=$? wall=${SECONDS}s"; tail -3 gpurun_out/r02_bench_n8_final.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n8_final.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'halo', d['config']['halo'][:20])
print('parity', d.get('parity',{}).get('nccl_rel_l2'), [c.get('halo_rank0') for c in d['parity']['cases']])
print('ab', d.get('halo_comparison'))
print('vcycle', d.get('vcycle',{}).get('seconds_per_vcycle'), d.get('vcycle',{}).get('residual_reduction_per_vcycle'))
a=d.get('affine_fast_path',{}); print('affine', a.get('GDOF_per_s'), a.get('vcycle',{}).get('seconds_per_vcycle'))
print('sphere', {k:round(v['GDOF_per_s'],2) for k,v in d['cubed_sphere'].items() if isinstance(v,dict)})
print('punct', d['two_punctures']['jacobian_apply']['GDOF_per_s'])
PY
