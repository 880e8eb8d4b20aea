#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/t_parity.log 2>&1; echo "parity rc=$?"
tail -4 gpurun_out/t_parity.log
python profiles/decode_probe.py llama-3.2-1b 6 > gpurun_out/probe_plain.log 2>&1 &&
RT_PROFILE_DECODE=1 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r1a.csv python profiles/decode_probe.py llama-3.2-1b 6 > gpurun_out/ncu_launch.log 2>&1
echo "ncu launch rc=$?"; tail -2 gpurun_out/probe_plain.log
RT_PROFILE_DECODE=1 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemv_tc -c 6 -o gpurun_out/gemv_r1a python profiles/decode_probe.py llama-3.2-1b 3 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
