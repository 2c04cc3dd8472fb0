cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_smoother.py tests/test_multigrid.py tests/test_krylov.py tests/test_gpu_apply.py -m gpu -q -x 2>&1 | tail -25) > gpurun_out/r02_fused_tests.txt
cat gpurun_out/r02_fused_tests.txt
for f in 1 0; do
D4B_FUSED_SOLVERS=$f timeout 600 python bench.py --steps 20 --warmup 5 --no-sphere --no-punctures --no-sweep --no-cpu > gpurun_out/r02_bench_fused$f.json 2> gpurun_out/r02_bench_fused$f.err
echo "fused=$f rc=$?"; tail -3 gpurun_out/r02_bench_fused$f.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_fused$f.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'])
print('vcycle', d.get('vcycle',{}).get('seconds_per_vcycle'), d.get('vcycle',{}).get('residual_reduction_per_vcycle'))
a=d.get('affine_fast_path',{}); print('affine', a.get('GDOF_per_s'), a.get('vcycle',{}).get('seconds_per_vcycle'))
PY
done
