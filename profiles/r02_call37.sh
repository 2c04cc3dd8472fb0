cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_smoother.py -m gpu -q -x -k "fused_solvers_on_bricks or match_reference" 2>&1 | tail -8
echo "memcheck rc=$?"
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_smoother.py -m gpu -q -x -k "fused_solvers_on_bricks and (3 or 2)" 2>&1 | tail -8
echo "racecheck rc=$?"
