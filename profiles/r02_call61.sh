#!/bin/bash
# hp mortar kernel: L1 prefetch of the neighbour's uh at the top of its face (pfnbr0) or one face ahead (pfnbr1)
mkdir -p gpurun_out
{
for lp in "5 3" "4 5"; do
for v in pfnbr0 pfnbr1; do
D4B_LIBRARY_PATH=profiles/variants/lib_$v.so timeout 300 python profiles/apply_hp_once.py $lp
done
timeout 300 python profiles/apply_hp_once.py $lp
done
} 2>&1 | grep random | tee gpurun_out/r02_hp_pfnbr.txt
