#!/bin/bash
cd theroundtaible_b200/_lib
for st in 2 3 4; do echo "== stages $st"; cp librt_stage$st.so librt_engine.so; (cd ../.. && python scripts/attn_bench.py 2>&1 | grep "attn bench"); done
