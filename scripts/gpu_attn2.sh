#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/t_par.log 2>&1; echo "parity rc=$?"; tail -4 gpurun_out/t_par.log
for v in 1 0; do echo -n "1B v2=$v "; RT_ATTN_V2=$v python profiles/decode_probe.py llama-3.2-1b 128 | tail -1; done
for v in 1 0; do echo -n "8B v2=$v "; RT_ATTN_V2=$v python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 | tail -1; done
for cp in 4 16; do echo -n "8B v2 chunk_pages=$cp "; RT_ATTN_MIN_CHUNK_PAGES=$cp python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 | tail -1; done
for cp in 4 8; do echo -n "1B v2 chunk_pages=$cp "; RT_ATTN_MIN_CHUNK_PAGES=$cp python profiles/decode_probe.py llama-3.2-1b 128 | tail -1; done
