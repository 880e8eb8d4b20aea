#!/bin/bash
# ncu launch lists (durations + DRAM bytes) of a few decode steps of tensor-parallel SHARDS run alone on one GPU (ncu cannot
# wrap a multi-rank command; the exchange kernels are in the in-kernel timelines instead).  Each command runs plain first.
mkdir -p gpurun_out
TAG=${1:-r2}
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread"
python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 4 > gpurun_out/plain_s1.log 2>&1 &&
RT_PROFILE_PHASE=decode ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/${TAG}_launches_decode_shard70b_tp8.csv --metrics $M \
  python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 4 > gpurun_out/ncu_s1.log 2>&1; echo "70b/tp8 shard rc=$?"
python scripts/tp_shard_probe.py llama-3.1-8b 8 5 8192 128 4 > gpurun_out/plain_s2.log 2>&1 &&
RT_PROFILE_PHASE=decode ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/${TAG}_launches_decode_shard8b_tp8.csv --metrics $M \
  python scripts/tp_shard_probe.py llama-3.1-8b 8 5 8192 128 4 > gpurun_out/ncu_s2.log 2>&1; echo "8b/tp8 shard rc=$?"
