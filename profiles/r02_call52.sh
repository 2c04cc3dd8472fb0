#!/bin/bash
# hp mortar kernel: lane groups (8/16/32 lanes per element) against one warp per element
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_gpu_apply.py tests/test_sphere.py tests/test_smoother.py tests/test_amr_brick.py tests/test_newton.py tests/test_dropin.py -m gpu -x -q 2>&1 | tail -4
for i in 1 2; do
D4B_LIBRARY_PATH=profiles/variants/lib_hpwarp.so timeout 300 python profiles/apply_hp_once.py 5 3
timeout 300 python profiles/apply_hp_once.py 5 3
done
D4B_LIBRARY_PATH=profiles/variants/lib_hpwarp.so timeout 300 python profiles/apply_hp_once.py 4 2
timeout 300 python profiles/apply_hp_once.py 4 2
} 2>&1 | tee gpurun_out/r02_hp_groups.txt
