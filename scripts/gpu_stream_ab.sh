#!/bin/bash
for s in 1 0; do echo -n "1B stream=$s "; RT_GEMV_STREAM=$s timeout -s KILL 200 python profiles/decode_probe.py llama-3.2-1b 128 | tail -1; done
for s in 1 0; do echo -n "8B stream=$s "; RT_GEMV_STREAM=$s timeout -s KILL 300 python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 | tail -1; done
