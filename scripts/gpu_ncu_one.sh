#!/bin/bash
# ONE profiler pass per GPU call, after the same command has exited 0 without the profiler.
# usage: gpu_ncu_one.sh list-decode | full-decode | list-prefill | full-prefill   [tag]
mode=$1; tag=${2:-r1c}
mkdir -p gpurun_out
case $mode in
  list-decode)  phase=decode;  n=6; out="--csv --log-file gpurun_out/${tag}_launches_decode_config2.csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,launch__registers_per_thread";;
  list-prefill) phase=prefill; n=2; out="--csv --log-file gpurun_out/${tag}_launches_prefill_config2.csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,launch__registers_per_thread";;
  full-decode)  phase=decode;  n=3; out="--set full --import-source on -k regex:gemv_kernel|attn_decode -c 10 -o gpurun_out/${tag}_decode_full";;
  full-prefill) phase=prefill; n=2; out="--set full --import-source on -k regex:gemm_tc5|attn_prefill -c 6 -o gpurun_out/${tag}_prefill_full";;
  *) echo "unknown mode"; exit 2;;
esac
export RT_PROFILE_PHASE=$phase
python profiles/decode_probe.py llama-3.2-1b $n > gpurun_out/probe_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/probe_plain.log; exit 1; }
tail -1 gpurun_out/probe_plain.log
ncu --profile-from-start off --clock-control none $out python profiles/decode_probe.py llama-3.2-1b $n > gpurun_out/ncu_$mode.log 2>&1
echo "ncu $mode rc=$?"; tail -2 gpurun_out/ncu_$mode.log
