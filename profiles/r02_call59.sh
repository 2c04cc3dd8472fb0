#!/bin/bash
# hp mortar kernel: (-)-side metric planes copied asynchronously into the flux buffer (variant asyncm) against the committed kernel
mkdir -p gpurun_out
{
for lp in "5 3" "4 5" "4 2"; do
D4B_LIBRARY_PATH=profiles/variants/lib_asyncm.so timeout 300 python profiles/apply_hp_once.py $lp
timeout 300 python profiles/apply_hp_once.py $lp
done
} 2>&1 | grep random | tee gpurun_out/r02_hp_asyncm.txt
