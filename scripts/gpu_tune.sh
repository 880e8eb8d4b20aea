#!/bin/bash
mkdir -p gpurun_out
for mt in 1 2 4; do for nw in 8 16; do RT_GEMV_MT=$mt RT_GEMV_NW=$nw python scripts/gemv_tune.py 3; done; done 2>&1 | tee gpurun_out/tune.log
RT_NO_PDL=1 python scripts/gemv_tune.py 3 | tee -a gpurun_out/tune.log
python scripts/gemv_tune.py 3 | tee -a gpurun_out/tune.log
bash scripts/gpu_iter.sh
