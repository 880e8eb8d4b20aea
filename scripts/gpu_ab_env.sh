#!/bin/bash
# A/B of an environment switch on the decode probes:   gpu_ab_env.sh VAR=val [VAR=val ...]   (default first, then with the switches)
run() {
  python profiles/decode_probe.py llama-3.1-8b 48 5 8192 128 2>&1 | tail -1
  python profiles/decode_probe.py llama-3.2-1b 128 2>&1 | tail -1
  python scripts/tp_shard_probe.py llama-3.1-70b 8 8 16384 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py llama-3.1-8b 8 5 8192 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py llama-3.1-8b 4 5 8192 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py llama-3.1-8b 2 5 8192 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py qwen2.5-32b 4 8 2048 128 32 2>&1 | tail -1
  python scripts/tp_shard_probe.py qwen2.5-32b 2 8 2048 128 32 2>&1 | tail -1
}
echo "== default"; run
echo "== $@"; env "$@" bash -c "$(declare -f run); run"
