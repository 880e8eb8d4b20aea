#!/bin/bash
bash scripts/gpu_iter.sh
for m in 0 4 8 16 32 64; do echo -n "skip=$m "; RT_SKIP_MASK=$m python profiles/decode_probe.py llama-3.2-1b 128 | tail -1; done
