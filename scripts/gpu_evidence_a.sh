#!/bin/bash
# evidence, part A (no profiler): full GPU tests, bench config 2 (with the CPU baseline leg and the discussion leg) and config 3
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_all.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_config2.json 2> gpurun_out/bench.err; echo "bench2 rc=$?"
python bench.py --workload config3 --steps 2 --warmup 3 > gpurun_out/bench_config3.json 2>> gpurun_out/bench.err; echo "bench3 rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2>> gpurun_out/bench.err; echo "ref rc=$?"
python - <<'PY'
import json
for f in ("bench_config2", "bench_config3", "bench_reference"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["value"], 1), d["unit"], "ms/step", round(d["ms_per_step"], 1), "frac", d.get("roofline", {}).get("frac"), "e2e", round(d["e2e"]["value"], 1))
    except Exception as ex:
        print(f, "parse failed", ex)
PY
tail -3 gpurun_out/bench.err
