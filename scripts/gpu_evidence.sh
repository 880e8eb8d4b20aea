#!/bin/bash
# round-1 evidence: tests, bench (config 2 and 3), ncu launch lists (decode + prefill), ncu --set full of the top kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_all.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_config2.json 2> gpurun_out/bench.err; echo "bench2 rc=$?"
python bench.py --workload config3 --steps 2 --warmup 1 > gpurun_out/bench_config3.json 2>> gpurun_out/bench.err; echo "bench3 rc=$?"
python profiles/decode_probe.py llama-3.2-1b 6 > gpurun_out/probe_plain.log 2>&1 &&
RT_PROFILE_PHASE=decode ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,launch__registers_per_thread --clock-control none --csv --log-file gpurun_out/r1b_launches_decode_config2.csv python profiles/decode_probe.py llama-3.2-1b 6 > gpurun_out/ncu1.log 2>&1
echo "ncu decode list rc=$?"
RT_PROFILE_PHASE=prefill ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,launch__registers_per_thread --clock-control none --csv --log-file gpurun_out/r1b_launches_prefill_config2.csv python profiles/decode_probe.py llama-3.2-1b 2 > gpurun_out/ncu2.log 2>&1
echo "ncu prefill list rc=$?"
RT_PROFILE_PHASE=decode ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemv_kernel -c 6 -o gpurun_out/r1b_gemv python profiles/decode_probe.py llama-3.2-1b 3 > gpurun_out/ncu3.log 2>&1
echo "ncu gemv full rc=$?"
RT_PROFILE_PHASE=prefill ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"gemm_tc5|attn_prefill" -c 6 -o gpurun_out/r1b_prefill python profiles/decode_probe.py llama-3.2-1b 2 > gpurun_out/ncu4.log 2>&1
echo "ncu prefill full rc=$?"
