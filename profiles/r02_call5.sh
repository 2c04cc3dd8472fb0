cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_apply.py tests/test_sphere.py tests/test_smoother.py tests/test_newton.py -q -x 2>&1 | tail -8) > gpurun_out/r02_curved_tests.txt
cat gpurun_out/r02_curved_tests.txt
for v in 0 1; do
  D4B_CURVED_SHFL=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-vcycle --no-sweep --no-punctures > gpurun_out/r02_sphere_shfl$v.json 2> gpurun_out/r02_sphere_shfl$v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r02_sphere_shfl$v.json').read().strip().splitlines()[-1])
print('SHFL=$v', {k:(round(v['GDOF_per_s'],2), round(v.get('ms_per_apply',0),3)) for k,v in d['cubed_sphere'].items() if isinstance(v,dict)})
PY
done
