#!/bin/bash
# attention stage (tcgen05 decode kernel + merge): parity tests, shard / config-3 step times, in-kernel timelines
mkdir -p gpurun_out
TAG=${1:-r2c}
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/t_attn.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_attn.log
python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1
python profiles/decode_probe.py llama-3.2-1b 128 2>&1 | tail -1
python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 32 2>&1 | tail -1
python scripts/tp_shard_probe.py llama-3.1-8b 8 5 8192 128 32 2>&1 | tail -1
python scripts/tp_shard_probe.py llama-3.1-8b 4 5 8192 128 32 2>&1 | tail -1
python scripts/tp_shard_probe.py qwen2.5-32b 4 8 2048 128 32 2>&1 | tail -1
echo "RT_DEC5_EARLY=0"
RT_DEC5_EARLY=0 python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 32 2>&1 | tail -1
RT_DEC5_EARLY=0 python scripts/tp_shard_probe.py llama-3.1-8b 8 5 8192 128 32 2>&1 | tail -1
echo "RT_MERGE_WIDE=1000"
RT_MERGE_WIDE=1000 python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 32 2>&1 | tail -1
RT_MERGE_WIDE=1000 python scripts/tp_shard_probe.py llama-3.1-8b 8 5 8192 128 32 2>&1 | tail -1
RT_MERGE_WIDE=1000 python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1
bash scripts/gpu_trace_shard.sh $TAG
