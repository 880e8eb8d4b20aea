#!/bin/bash
# in-kernel timeline (trace build) of one layer of a tensor-parallel SHARD run alone on one GPU
#   usage: gpu_trace_shard.sh <tag>      (python -m theroundtaible_b200.build --trace first)
mkdir -p gpurun_out
TAG=${1:-r2}
export RT_LIB=$PWD/theroundtaible_b200/_lib/librt_engine_trace.so
RT_TRACE_FILE=gpurun_out/${TAG}_70b8.bin python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 8 2>&1 | tail -1
python scripts/trace_chain.py gpurun_out/${TAG}_70b8.bin 12 > gpurun_out/${TAG}_shard70b_tp8_chain.txt; cat gpurun_out/${TAG}_shard70b_tp8_chain.txt
RT_TRACE_FILE=gpurun_out/${TAG}_8b8.bin python scripts/tp_shard_probe.py llama-3.1-8b 8 5 8192 128 8 2>&1 | tail -1
python scripts/trace_chain.py gpurun_out/${TAG}_8b8.bin 12 > gpurun_out/${TAG}_shard8b_tp8_chain.txt; cat gpurun_out/${TAG}_shard8b_tp8_chain.txt
rm -f gpurun_out/${TAG}_70b8.bin gpurun_out/${TAG}_8b8.bin
