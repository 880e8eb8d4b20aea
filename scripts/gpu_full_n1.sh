#!/bin/bash
# whole GPU suite, then the default bench line (N = 1)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/t_all.log 2>&1; echo "all rc=$?"; tail -4 gpurun_out/t_all.log
timeout 1500 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_n1.json
