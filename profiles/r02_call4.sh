cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_newton.py tests/test_multigrid.py -q 2>&1 | tail -8) > gpurun_out/r02_newton_tests.txt
cat gpurun_out/r02_newton_tests.txt
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu --no-vcycle --no-sphere --no-sweep > gpurun_out/r02_bench_punct.json 2> gpurun_out/r02_bench_punct.err
echo rc=$?; tail -c 2500 gpurun_out/r02_bench_punct.json; tail -5 gpurun_out/r02_bench_punct.err
