#!/bin/bash
# hp mortar kernel: table offsets in constant memory (working tree of that call, not kept) against the pointer in the arguments (gtabs); hp tests
mkdir -p gpurun_out
{
for lp in "5 3" "4 5"; do
D4B_LIBRARY_PATH=profiles/variants/lib_gtabs.so timeout 300 python profiles/apply_hp_once.py $lp
timeout 300 python profiles/apply_hp_once.py $lp
done
} 2>&1 | grep random | tee gpurun_out/r02_hp_consttabs.txt
timeout 600 python -m pytest tests/test_gpu_apply.py tests/test_sphere.py tests/test_amr_brick.py -m gpu -x -q 2>&1 | tail -3
