#!/bin/bash
# fp16-twin decode mode (gemv.cuh mode 2) against the bf16 hi + lo mode: parity tests, then config-3 / shard step times both ways
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_kernels.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/t_f16.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/t_f16.log
for f in 1 0; do
  echo "RT_DEC_F16=$f"
  RT_DEC_F16=$f python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1
  RT_DEC_F16=$f python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 32 2>&1 | tail -1
  RT_DEC_F16=$f python scripts/tp_shard_probe.py qwen2.5-32b 4 8 2048 128 32 2>&1 | tail -1
done
