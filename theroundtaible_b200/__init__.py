"""B200-native generation backend for TheRoundtAIble's `local-llm` adapter.

Layout: ``csrc/`` (CUDA kernels for sm_100a + the C-ABI of include/rt_engine.h), ``engine.py`` (ctypes
binding of that ABI), ``adapter.py`` (mirror of the reference's adapter interface).  No CPU fallback.
"""
from .adapter import AdapterError, B200LocalLlmAdapter, BaseAdapter, classify_error, create_adapter  # noqa: F401
from .engine import Engine, EngineError  # noqa: F401

__all__ = ["AdapterError", "B200LocalLlmAdapter", "BaseAdapter", "classify_error", "create_adapter", "Engine",
           "EngineError"]
