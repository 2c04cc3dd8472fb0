cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_dropin.py -q -x 2>&1 | tail -30) > gpurun_out/r02_dropin.txt
cat gpurun_out/r02_dropin.txt
(timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15) > gpurun_out/r02_gpu_suite_1gpu.txt
cat gpurun_out/r02_gpu_suite_1gpu.txt
