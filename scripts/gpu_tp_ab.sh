#!/bin/bash
# usage: gpu_tp_ab.sh <ngpu> <workload> <layers>
N=$1; W=$2; L=$3
mkdir -p gpurun_out
for f in 1 0; do
RT_TP_FUSED=$f timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$f bench.py --gpus $N --tp $N --workload $W --layers $L --steps 2 --warmup 1 > gpurun_out/ab_${W}_tp${N}_f$f.json 2> gpurun_out/ab_f$f.err; echo "fused=$f rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/ab_${W}_tp${N}_f$f.json').read().strip().splitlines()[-1]); print("fused=$f value", round(d["value"]), "step_ms", round(d["config"]["decode_ms_per_token_step"],3), "frac", round(d["roofline"]["frac"],3), "prefill_ms", round(d["prefill"]["ms"],1))
except Exception as ex: print("parse fail", ex)
PY
done
grep -v "OMP_NUM\|\*\*\*\|^$" gpurun_out/ab_f1.err | tail -3
