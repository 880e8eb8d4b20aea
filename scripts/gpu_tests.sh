#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/t_all.log 2>&1; echo "all rc=$?"; tail -6 gpurun_out/t_all.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/t_zz.log 2>&1; echo "zz alone rc=$?"; tail -4 gpurun_out/t_zz.log
