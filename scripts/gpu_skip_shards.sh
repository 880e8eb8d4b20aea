#!/bin/bash
# marginal cost of each kernel class in the dependent chain of a decode step (RT_SKIP_MASK leaves the class out; results
# are then wrong, timing only): 2 qkv, 4 attention (+ merge), 8 o-proj, 16 gate/up, 32 down, 64 lm_head
for cfg in "llama-3.1-70b 8 8 16384 128 24" "llama-3.1-8b 8 5 8192 128 24" "qwen2.5-32b 4 8 2048 128 24"; do
  for m in 0 2 4 8 16 32 64; do echo -n "skip=$m "; RT_SKIP_MASK=$m python scripts/tp_shard_probe.py $cfg 2>&1 | tail -1 | cut -c1-90; done
done
