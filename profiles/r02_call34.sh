cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python profiles/apply_sphere_once.py 5 3 | tail -1
ncu --set full --clock-control none --import-source on -k regex:curved_apply_kernel -s 2 -c 1 -o gpurun_out/r02_curved_p3_sphere -f python profiles/apply_sphere_once.py 5 3 > gpurun_out/ncu_curved.log 2>&1
tail -2 gpurun_out/ncu_curved.log
