#!/bin/bash
# full GPU suite on the final source of the round
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r02_gpu_suite_final.txt
