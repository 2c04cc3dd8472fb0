cd $GRAFT_REPO_ROOT
(timeout 1200 python -m pytest tests/test_newton.py tests/test_dropin.py tests/test_quadrature_ops.py -m gpu -q -x 2>&1 | tail -4)
for f in 1 0; do
D4B_FUSED_RESIDUAL=$f python bench.py --steps 10 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-sweep > gpurun_out/b_res$f.json 2> gpurun_out/b_res$f.err; tail -1 gpurun_out/b_res$f.err
python - <<PY
import json
d=json.loads(open('gpurun_out/b_res$f.json').read().strip().splitlines()[-1])
t=d['two_punctures']; n=t['newton_solve']; print('fused=$f', 'jac', round(t['jacobian_apply']['GDOF_per_s'],2), 'res', round(t['residual']['GDOF_per_s'],2), round(t['residual']['ms_per_evaluation'],3), 'newton', round(n['seconds'],3), n['fcg_iterations_per_step'], n['residual_norms'][-1])
PY
done
