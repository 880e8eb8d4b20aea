#!/bin/bash
# sweep of the L2 weight-prefetch budgets (RT_PF_ATTN_MB: during attention; RT_PF_TAIL_MB: at each projection's tail)
for cfg in "0 0" "16 0" "32 0" "48 0" "64 0" "0 2" "0 4" "0 8" "32 4" "48 4" "48 8"; do
  set -- $cfg
  echo -n "1B attn=$1 tail=$2 "; RT_PF_ATTN_MB=$1 RT_PF_TAIL_MB=$2 python profiles/decode_probe.py llama-3.2-1b 128 | tail -1
done
for cfg in "0 0" "32 0" "64 0" "96 0" "0 8" "64 8" "96 16"; do
  set -- $cfg
  echo -n "8B attn=$1 tail=$2 "; RT_PF_ATTN_MB=$1 RT_PF_TAIL_MB=$2 python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 | tail -1
done
