#!/bin/bash
# ncu launch lists (durations + DRAM bytes) of: 3 decode steps of config 3, 5 decode steps of config 2, the prefill of config 3.
# Each profiled command is first run plain (must exit 0).  usage: gpu_ncu_lists.sh [tag]
mkdir -p gpurun_out
TAG=${1:-r2}
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread"
python profiles/decode_probe.py llama-3.1-8b 4 5 8192 128 > gpurun_out/plain1.log 2>&1 &&
RT_PROFILE_PHASE=decode ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/${TAG}_launches_decode_config3.csv --metrics $M \
  python profiles/decode_probe.py llama-3.1-8b 4 5 8192 128 > gpurun_out/ncu1.log 2>&1; echo "config3 decode rc=$?"
python profiles/decode_probe.py llama-3.2-1b 6 > gpurun_out/plain2.log 2>&1 &&
RT_PROFILE_PHASE=decode ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/${TAG}_launches_decode_config2.csv --metrics $M \
  python profiles/decode_probe.py llama-3.2-1b 6 > gpurun_out/ncu2.log 2>&1; echo "config2 decode rc=$?"
python profiles/decode_probe.py llama-3.1-8b 2 5 8192 128 > gpurun_out/plain3.log 2>&1 &&
RT_PROFILE_PHASE=prefill ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/${TAG}_launches_prefill_config3.csv --metrics gpu__time_duration.sum \
  python profiles/decode_probe.py llama-3.1-8b 2 5 8192 128 > gpurun_out/ncu3.log 2>&1; echo "config3 prefill rc=$?"
