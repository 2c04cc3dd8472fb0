#!/bin/bash
# hp mortar kernel: L2 prefetch of all six neighbours' uh and mortar metric at descriptor time (variant l2pf) against
# the committed kernel; ncu capture of the committed kernel
mkdir -p gpurun_out
{
for lp in "5 3" "4 5"; do
D4B_LIBRARY_PATH=profiles/variants/lib_l2pf.so timeout 300 python profiles/apply_hp_once.py $lp
timeout 300 python profiles/apply_hp_once.py $lp
done
} 2>&1 | tee gpurun_out/r02_hp_l2pf.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:hp_apply_kernel -s 3 -c 3 -o gpurun_out/r02_hp_kernel_after -f python profiles/apply_hp_once.py 4 3 > gpurun_out/r02_hp_ncu2.log 2>&1
tail -2 gpurun_out/r02_hp_ncu2.log
