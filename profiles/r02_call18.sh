cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -15) > gpurun_out/r02_gpu_suite_mid.txt
cat gpurun_out/r02_gpu_suite_mid.txt
python -c "import __graft_entry__ as g; g.smoke()"
python profiles/apply_time.py 6:3 6:1 6:2 6:4 6:5 6:6 6:7 2>&1 | grep level | tee gpurun_out/r02_apply_times_mid.txt
