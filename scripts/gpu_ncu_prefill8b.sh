#!/bin/bash
# launch list of the prefill phase of config 3 (8B, 8,192 shared + 5 x 128 own tokens): one profiler pass
# usage: gpu_ncu_prefill8b.sh [tc5]
mkdir -p gpurun_out
export RT_PROFILE_PHASE=prefill
if [ "$1" = "tc5" ]; then export RT_PFA_TC5=1; fi
python profiles/decode_probe.py llama-3.1-8b 2 5 8192 128 > gpurun_out/probe_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
tail -1 gpurun_out/probe_plain.log
ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/r1c_launches_prefill_config3_$1.csv --metrics gpu__time_duration.sum,launch__registers_per_thread python profiles/decode_probe.py llama-3.1-8b 2 5 8192 128 > gpurun_out/ncu_pf8b.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/ncu_pf8b.log
