#!/bin/bash
# A/B of two builds of the library on the decode probes: default build, then theroundtaible_b200/_lib/librt_engine_<tag>.so
TAGS="${@:-mb2}"
run() {
  python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1
  python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py llama-3.1-8b 8 5 8192 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py llama-3.1-8b 4 5 8192 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py llama-3.1-8b 2 5 8192 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py qwen2.5-32b 4 8 2048 128 32 2>&1 | tail -1
}
echo "== default"; run
for t in $TAGS; do echo "== $t"; RT_LIB=$PWD/theroundtaible_b200/_lib/librt_engine_$t.so run; done
