cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python profiles/apply_once.py 6 3
ncu --set full --clock-control none --import-source on -k regex:flat_apply_kernel -s 2 -c 1 -o gpurun_out/r02_flat_p3 -f python profiles/apply_once.py 6 3 > gpurun_out/ncu_flat.log 2>&1
tail -3 gpurun_out/ncu_flat.log
ls -la gpurun_out/*.ncu-rep
