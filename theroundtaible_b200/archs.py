"""Model-name table of the backend (host side).  The reference carries only a model NAME
(/root/reference/src/types.ts:29-34 AdapterLocalConfig.model); the native library resolves it to
dimensions (rt_arch_by_name).  This module lists the names and a few derived sizes used by the
adapter and the benchmark."""
from __future__ import annotations

MODELS = ("qwen2.5-0.5b", "llama-3.2-1b", "llama-3.1-8b", "qwen2.5-32b", "llama-3.1-70b",
          "tiny-llama", "tiny-llama-hd128", "tiny-qwen", "tiny-qwen-hd128")
