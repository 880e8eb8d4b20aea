#!/bin/bash
for m in 0 2 4 8 16 32 64; do echo -n "skip=$m "; RT_SKIP_MASK=$m python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 | tail -1; done
