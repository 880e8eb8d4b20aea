#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 120 python -m pytest tests/test_gpu_kernels.py -m gpu -q --tb=short -p no:cacheprovider -x -k "gemv" > gpurun_out/t_gemv.log 2>&1; echo "gemv tests rc=$?"; tail -8 gpurun_out/t_gemv.log
timeout -s KILL 120 python scripts/gemv_tune.py 3 2>&1 | tail -2
RT_GEMV_STREAM=0 timeout -s KILL 120 python scripts/gemv_tune.py 3 2>&1 | tail -1
