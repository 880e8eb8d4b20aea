#!/bin/bash
mkdir -p gpurun_out
python profiles/decode_probe.py llama-3.2-1b 6 > gpurun_out/probe_plain.log 2>&1 &&
RT_PROFILE_DECODE=1 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,launch__registers_per_thread --clock-control none --csv --log-file gpurun_out/launches.csv python profiles/decode_probe.py llama-3.2-1b 6 > gpurun_out/ncu_launch.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/probe_plain.log
RT_NO_PDL=1 python profiles/decode_probe.py llama-3.2-1b 64 | tail -1
python profiles/decode_probe.py llama-3.2-1b 64 | tail -1
