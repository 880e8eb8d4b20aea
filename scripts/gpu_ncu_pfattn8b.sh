#!/bin/bash
# one full-set capture of the prefill attention kernel on the config-3 shape (8B, head_dim 128, 8k keys)
# usage: gpu_ncu_pfattn8b.sh [tc5]   (tc5 = the tcgen05 kernel, RT_PFA_TC5=1)
mkdir -p gpurun_out
export RT_PROFILE_PHASE=prefill
if [ "$1" = "tc5" ]; then export RT_PFA_TC5=1; fi
python profiles/decode_probe.py llama-3.1-8b 2 5 8192 128 > gpurun_out/probe_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
tail -1 gpurun_out/probe_plain.log
ncu --profile-from-start off --clock-control none --set full --import-source on -k regex:attn_prefill -s 100 -c 2 -o gpurun_out/r1c_pfattn8b_$1 python profiles/decode_probe.py llama-3.1-8b 2 5 8192 128 > gpurun_out/ncu_pfattn.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/ncu_pfattn.log
