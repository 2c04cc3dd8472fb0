cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -15) > gpurun_out/r02_gpu_suite_devsetup.txt
cat gpurun_out/r02_gpu_suite_devsetup.txt
D4B_TIMING=1 python profiles/hp_bench.py 5 3 uniform 2>&1 | tail -12
timeout 600 ncu --set full --clock-control none --import-source on -k regex:curved_apply_kernel -c 1 -o gpurun_out/r02_curved_p3 python profiles/hp_bench.py 5 3 uniform > gpurun_out/r02_ncu_curved.log 2>&1
tail -3 gpurun_out/r02_ncu_curved.log
ls -la gpurun_out/*.ncu-rep
