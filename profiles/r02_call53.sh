#!/bin/bash
# hp mortar kernel: lane-group sizes at N=3,4 and one ncu capture of the kernel
mkdir -p gpurun_out
{
for v in hpwarp gs4_16 gs3_8_gs4_16; do
D4B_LIBRARY_PATH=profiles/variants/lib_$v.so timeout 300 python profiles/apply_hp_once.py 5 3
done
timeout 300 python profiles/apply_hp_once.py 5 3
for v in hpwarp gs4_16; do
D4B_LIBRARY_PATH=profiles/variants/lib_$v.so timeout 300 python profiles/apply_hp_once.py 4 5
done
timeout 300 python profiles/apply_hp_once.py 4 5
} 2>&1 | tee gpurun_out/r02_hp_groups2.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:hp_apply_kernel -s 3 -c 3 -o gpurun_out/r02_hp_kernel -f python profiles/apply_hp_once.py 4 3 > gpurun_out/r02_hp_ncu.log 2>&1
tail -3 gpurun_out/r02_hp_ncu.log
