#!/bin/bash
# 2-GPU confirmation: the tensor-parallel tests, then the bench line at N = 2 (strong scaling, TP = 2)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/t_tp2.log 2>&1; echo "tp tests rc=$?"; tail -4 gpurun_out/t_tp2.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_n2.json
