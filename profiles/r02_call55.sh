#!/bin/bash
# hp mortar kernel: lane-group size chosen by the launcher from the shared-memory footprint; against one warp per element
mkdir -p gpurun_out
{
for lp in "5 3" "4 5" "4 2" "4 7"; do
D4B_LIBRARY_PATH=profiles/variants/lib_hpwarp.so timeout 300 python profiles/apply_hp_once.py $lp
timeout 300 python profiles/apply_hp_once.py $lp
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
} 2>&1 | tee gpurun_out/r02_hp_groups3.txt
