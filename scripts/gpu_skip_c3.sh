#!/bin/bash
# config 3 on one GPU: marginal cost of each kernel class (RT_SKIP_MASK), the tcgen05 decode attention forced, merge widths
for m in 0 2 4 8 16 32 64; do echo -n "skip=$m "; RT_SKIP_MASK=$m python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1; done
echo -n "dec5 forced "; RT_DEC5=2 python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1
echo -n "dec5 forced, wide merge "; RT_DEC5=2 RT_MERGE_WIDE_CTAS=1000 python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1
echo -n "8b tp2 shard, narrow merge "; RT_MERGE_WIDE_CTAS=0 python scripts/tp_shard_probe.py llama-3.1-8b 2 5 8192 128 32 2>&1 | tail -1
echo -n "8b tp2 shard "; python scripts/tp_shard_probe.py llama-3.1-8b 2 5 8192 128 32 2>&1 | tail -1
echo -n "32b tp4 shard, narrow merge "; RT_MERGE_WIDE_CTAS=0 python scripts/tp_shard_probe.py qwen2.5-32b 4 8 2048 128 32 2>&1 | tail -1
echo -n "32b tp4 shard "; python scripts/tp_shard_probe.py qwen2.5-32b 4 8 2048 128 32 2>&1 | tail -1
