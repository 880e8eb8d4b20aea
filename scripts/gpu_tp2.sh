#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/t_tp.log 2>&1; echo "tp tests rc=$?"; tail -15 gpurun_out/t_tp.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --tp 2 --workload config4 --layers 4 --steps 2 --warmup 1 > gpurun_out/bench_tp2.log 2> gpurun_out/bench_tp2.err; echo "bench tp2 rc=$?"; tail -c 1500 gpurun_out/bench_tp2.log; tail -5 gpurun_out/bench_tp2.err
