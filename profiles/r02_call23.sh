cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
out=gpurun_out/r02_ns_linear_ab.txt; : > $out
export D4B_LIBRARY_PATH=$PWD/profiles/variants/lib_r02start.so
python profiles/apply_time.py 6:3 6:1 2>&1 | grep level | tee -a $out
unset D4B_LIBRARY_PATH
for pf in 2 3 1; do
echo "D4B_PREFETCH=$pf" | tee -a $out
D4B_PREFETCH=$pf python profiles/apply_time.py 6:3 6:1 6:2 6:4 5:5 5:6 5:7 2>&1 | grep level | tee -a $out
done
(timeout 1200 python -m pytest tests/test_gpu_apply.py tests/test_smoother.py -m gpu -q -x 2>&1 | tail -3)
