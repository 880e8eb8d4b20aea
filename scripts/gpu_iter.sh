#!/bin/bash
# iteration loop: full GPU tests, then the bench (short)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"; tail -12 gpurun_out/t_all.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
    print("value", round(d["value"]), "tok/s  step_ms", round(d["config"]["decode_ms_per_token_step"],4), "frac", round(d["roofline"]["frac"],3),
          "prefill_ms", round(d["config"]["prefill_ms_per_round"],1), "gu_frac", round(d["roofline"]["dominant_kernel"]["frac"],3),
          "e2e", round(d["e2e"]["value"]), "seq", round(d["extra"]["seq"]["decode_tok_per_s"]), "launches", d["gpu_launches"])
except Exception as ex:
    print("bench parse failed", ex)
PY
tail -3 gpurun_out/bench.err
