#!/bin/bash
# config-3 decode: timing probe, then an ncu launch list of 3 decode steps (durations + DRAM bytes read and written)
mkdir -p gpurun_out
TAG=${1:-r2}
python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1
python profiles/decode_probe.py llama-3.2-1b 128 2>&1 | tail -1
RT_PROFILE_PHASE=decode ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/${TAG}_launches_decode_config3.csv \
  --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread \
  python profiles/decode_probe.py llama-3.1-8b 4 5 8192 128 > gpurun_out/ncu_dec8b.log 2>&1
echo "ncu rc=$?"; tail -1 gpurun_out/ncu_dec8b.log
