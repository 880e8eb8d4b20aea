#!/bin/bash
RT_GEMV_STREAM=1 timeout -s KILL 120 python -m pytest tests/test_gpu_kernels.py -m gpu -q --tb=short -p no:cacheprovider -x -k "gemv" 2>&1 | tail -2
RT_GEMV_STREAM=1 timeout -s KILL 120 python scripts/gemv_tune.py 3 2>&1 | tail -1
RT_GEMV_STREAM=0 timeout -s KILL 120 python scripts/gemv_tune.py 3 2>&1 | tail -1
for s in 1 0; do echo -n "1B stream=$s "; RT_GEMV_STREAM=$s timeout -s KILL 200 python profiles/decode_probe.py llama-3.2-1b 128 | tail -1; done
for s in 1 0; do echo -n "8B stream=$s "; RT_GEMV_STREAM=$s timeout -s KILL 300 python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 | tail -1; done
