#!/bin/bash
# launch list of the decode phase of config 3 (8B, 5 knights, 8,192 shared + 128 own tokens): one profiler pass
mkdir -p gpurun_out
export RT_PROFILE_PHASE=decode
python profiles/decode_probe.py llama-3.1-8b 4 5 8192 128 > gpurun_out/probe_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
tail -1 gpurun_out/probe_plain.log
ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/r1c_launches_decode_config3.csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,launch__registers_per_thread python profiles/decode_probe.py llama-3.1-8b 4 5 8192 128 > gpurun_out/ncu_dec8b.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/ncu_dec8b.log
