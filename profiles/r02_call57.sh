#!/bin/bash
# hp mortar kernel: each lane's first mortar point's metric loaded into registers ahead of the stage (with and without the
# L1 prefetch) against the committed kernel
mkdir -p gpurun_out
{
for lp in "5 3" "4 5"; do
for v in regm regm_nopf; do
D4B_LIBRARY_PATH=profiles/variants/lib_$v.so timeout 300 python profiles/apply_hp_once.py $lp
done
timeout 300 python profiles/apply_hp_once.py $lp
done
} 2>&1 | grep random | tee gpurun_out/r02_hp_regmetric.txt
