#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests/test_gpu_tp.py -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/t_tp.log 2>&1; echo "tp tests (fused) rc=$?"; tail -6 gpurun_out/t_tp.log
for f in 1 0; do
RT_TP_FUSED=$f timeout -s KILL 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$f bench.py --gpus 2 --tp 2 --workload config4 --layers 8 --steps 2 --warmup 1 > gpurun_out/b_f$f.json 2> gpurun_out/b_f$f.err; echo "fused=$f rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/b_f$f.json').read().strip().splitlines()[-1]); print("fused=$f value", round(d["value"]), "step_ms", round(d["config"]["decode_ms_per_token_step"],3), "frac", round(d["roofline"]["frac"],3))
except Exception as ex: print("parse fail", ex)
PY
done
grep -v "OMP_NUM\|\*\*\*\|^$" gpurun_out/b_f1.err | tail -4
