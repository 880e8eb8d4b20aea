#!/bin/bash
# usage: gpu_tp_big.sh <ngpu> <workload> [layers]
N=$1; W=$2; L=${3:-0}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --tp $N --workload $W --layers $L --steps 2 --warmup 1 > gpurun_out/bench_${W}_tp$N.json 2> gpurun_out/bench_${W}_tp$N.err; echo "rc=$?"
tail -c 2500 gpurun_out/bench_${W}_tp$N.json; grep -v "OMP_NUM_THREADS\|\*\*\*\*\|^$" gpurun_out/bench_${W}_tp$N.err | tail -5
