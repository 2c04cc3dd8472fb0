cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for pol in 1 2; do
D4B_PEER_HALO_DEFAULT=$pol timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 --no-sphere --no-punctures --no-affine --no-vcycle > gpurun_out/r02_bench_n2_order$pol.json 2> gpurun_out/r02_bench_n2_order$pol.err
echo "policy=$pol rc=$?"; tail -1 gpurun_out/r02_bench_n2_order$pol.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n2_order$pol.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'])
print('ab', {k:v for k,v in d.get('halo_comparison').items() if k!='policy'})
PY
done
