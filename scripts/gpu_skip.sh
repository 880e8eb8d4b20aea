#!/bin/bash
for m in 0 2 4 8 16 32 64 62; do echo -n "skip=$m "; RT_SKIP_MASK=$m python profiles/decode_probe.py llama-3.2-1b 128 | tail -1; done
echo -n "nopdl "; RT_NO_PDL=1 python profiles/decode_probe.py llama-3.2-1b 128 | tail -1
echo -n "nograph "; RT_NO_GRAPH=1 python profiles/decode_probe.py llama-3.2-1b 128 | tail -1
