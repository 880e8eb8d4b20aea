#!/bin/bash
# push form of the tensor-parallel exchange against the pull form, on N GPUs (default 2): tests, then step times both ways
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/t_tp_push.log 2>&1; echo "tp tests rc=$?"; tail -6 gpurun_out/t_tp_push.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for push in 1 0; do
  echo "== RT_TP_PUSH=$push"
  RT_TP_PUSH=$push timeout 600 $TR --master-port 29621 scripts/tp_probe.py llama-3.1-8b 5 8192 128 48 2>&1 | grep -a "us/step"
  RT_TP_PUSH=$push timeout 600 $TR --master-port 29622 scripts/tp_probe.py qwen2.5-32b 8 2048 128 32 2>&1 | grep -a "us/step"
done
# in-kernel timelines of rank 0 (trace build), push and pull
if [ -f theroundtaible_b200/_lib/librt_engine_trace.so ]; then
  for push in 1 0; do
    RT_TP_PUSH=$push RT_LIB=$PWD/theroundtaible_b200/_lib/librt_engine_trace.so timeout 600 $TR --master-port 29623 scripts/tp_probe.py llama-3.1-8b 5 8192 128 8 push$push 2>&1 | grep -a "us/step"
    python scripts/trace_chain.py gpurun_out/push${push}_trace_rank0.bin 14 > gpurun_out/r2d_tp${N}_config3_push${push}_chain.txt; cat gpurun_out/r2d_tp${N}_config3_push${push}_chain.txt
    rm -f gpurun_out/push${push}_trace_rank0.bin
  done
fi
