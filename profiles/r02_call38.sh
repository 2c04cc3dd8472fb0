cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests/test_dropin.py tests/test_krylov.py -m gpu -q -x 2>&1 | tail -4)
python bench.py --no-vcycle --no-sphere --no-affine --no-punctures --no-sweep > gpurun_out/b_plugin.json 2> gpurun_out/b_plugin.err; tail -2 gpurun_out/b_plugin.err
python - <<PY
import json
d=json.loads(open('gpurun_out/b_plugin.json').read().strip().splitlines()[-1])
print(json.dumps(d['e2e_apply_lhs'], indent=1))
print('value', d['value'], 'e2e', d['e2e']['value'])
PY
python profiles/apply_sphere_once.py 5 3 | tail -1
