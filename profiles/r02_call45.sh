cd $GRAFT_REPO_ROOT
(timeout 900 python -m pytest tests/test_multigrid.py tests/test_newton.py tests/test_dropin.py -m gpu -q -x 2>&1 | tail -3)
D4B_MG_DEBUG=1 python bench.py --steps 5 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-affine --no-sweep > gpurun_out/b_newton.json 2> gpurun_out/b_newton.err
grep "d4b mg" gpurun_out/b_newton.err | awk '{print $NF, $(NF-1), $(NF-2), $(NF-3), $(NF-4), $(NF-5)}' | sort | uniq -c | sort -rn | head
python - <<PY
import json
d=json.loads(open('gpurun_out/b_newton.json').read().strip().splitlines()[-1])
n=d['two_punctures']['newton_solve']; print(n['seconds'], n['fcg_iterations_per_step'], n['newton_steps'], n['residual_norms'][-1])
PY
