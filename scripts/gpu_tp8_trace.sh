#!/bin/bash
# in-kernel timelines of rank 0 at TP = N (trace build): config 5 (Llama-3.1-70B, 8 knights, 16k prefix) and the headline model
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
export RT_LIB=$PWD/theroundtaible_b200/_lib/librt_engine_trace.so
timeout 900 $TR --master-port 29641 scripts/tp_probe.py llama-3.1-70b 8 16384 128 8 c5 2>&1 | grep -a "us/step"
python scripts/trace_chain.py gpurun_out/c5_trace_rank0.bin 14 > gpurun_out/r2_tp${N}_config5_push_chain.txt; cat gpurun_out/r2_tp${N}_config5_push_chain.txt
timeout 600 $TR --master-port 29642 scripts/tp_probe.py llama-3.1-8b 5 8192 128 8 c3 2>&1 | grep -a "us/step"
python scripts/trace_chain.py gpurun_out/c3_trace_rank0.bin 14 > gpurun_out/r2_tp${N}_config3_push_chain.txt; cat gpurun_out/r2_tp${N}_config3_push_chain.txt
rm -f gpurun_out/c5_trace_rank0.bin gpurun_out/c3_trace_rank0.bin
