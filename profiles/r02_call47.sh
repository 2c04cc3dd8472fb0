cd $GRAFT_REPO_ROOT
python profiles/newton_vcycle_breakdown.py 4 2>&1 | tail -6
D4B_MG_DENSE_COARSE=0 python profiles/newton_vcycle_breakdown.py 4 2>&1 | tail -6 | head -4
(timeout 1200 python -m pytest tests/test_newton.py tests/test_multigrid.py tests/test_dropin.py -m gpu -q -x 2>&1 | tail -4)
python bench.py --steps 5 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-sweep > gpurun_out/b_newton.json 2> gpurun_out/b_newton.err; tail -1 gpurun_out/b_newton.err
python - <<PY
import json
d=json.loads(open('gpurun_out/b_newton.json').read().strip().splitlines()[-1])
n=d['two_punctures']['newton_solve']; print(n['seconds'], n['fcg_iterations_per_step'], n['newton_steps'], n['residual_norms'])
PY
