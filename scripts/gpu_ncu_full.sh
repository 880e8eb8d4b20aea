#!/bin/bash
# one `ncu --set full` capture of the dominant decode kernel (gate/up projection with the fused SwiGLU epilogue, config 3)
# after the same command has exited 0 without the profiler.  usage: gpu_ncu_full.sh [tag]
mkdir -p gpurun_out
TAG=${1:-r2}
python profiles/decode_probe.py llama-3.1-8b 4 5 8192 128 > gpurun_out/plain_full.log 2>&1 &&
RT_PROFILE_PHASE=decode ncu --profile-from-start off --set full --clock-control none --import-source on --kernel-name-base demangled \
  -k "regex:gemv_kernel<.int.1, .int.2, .int.2, .int.8, .int.2" -s 8 -c 2 -o gpurun_out/${TAG}_gemv_swiglu_full \
  python profiles/decode_probe.py llama-3.1-8b 4 5 8192 128 > gpurun_out/ncu_full.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/ncu_full.log
