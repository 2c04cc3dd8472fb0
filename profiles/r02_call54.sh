#!/bin/bash
# hp mortar kernel: face descriptors resolved by six lanes at once, mortar metric prefetched; each against the build without it
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_gpu_apply.py tests/test_sphere.py tests/test_amr_brick.py tests/test_smoother.py tests/test_dropin.py -m gpu -x -q 2>&1 | tail -4
for v in neither nodesc noprefetch; do
D4B_LIBRARY_PATH=profiles/variants/lib_$v.so timeout 300 python profiles/apply_hp_once.py 5 3
done
timeout 300 python profiles/apply_hp_once.py 5 3
D4B_LIBRARY_PATH=profiles/variants/lib_neither.so timeout 300 python profiles/apply_hp_once.py 4 5
timeout 300 python profiles/apply_hp_once.py 4 5
} 2>&1 | tee gpurun_out/r02_hp_desc_prefetch.txt
