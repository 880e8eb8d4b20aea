"""Seeded, counter-based weight generator (TEST INFRASTRUCTURE; see oracle/__init__.py).

Random-init weights "of the named architecture" (BASELINE.json north_star) must be the SAME
bytes in the CPU oracle and in the GPU engine, and a 70B model cannot be shipped as a file.
So both sides evaluate one pure function of (seed, tensor id, element index) built from
integer arithmetic plus ONE correctly-rounded fp32 multiply (and one fp32 add for norm
weights) followed by round-to-nearest-even to bf16 -- bit-exact on CPU (numpy) and GPU
(theroundtaible_b200/csrc/weights.cuh restates it with __fmul_rn/__fadd_rn).

Distribution: Irwin-Hall(4) over 16-bit uniforms, centred and scaled to std 0.02 (HF
``initializer_range``), i.e. a bell-shaped approximation of N(0, 0.02); norm weights are
1 + N(0, 0.02)-like, biases are N(0, 0.02)-like.  No counterpart in the reference (it holds
no weights: /root/reference/src/adapters/local-llm.ts:159 only names a model).
"""
from __future__ import annotations

import numpy as np

M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
IH4_STD = 37837.22722  # sqrt(4 * (65536**2 - 1) / 12)

# tensor kinds (tensor id = kind for globals, (layer+1)*16 + kind for per-layer tensors)
EMBED, LM_HEAD, FINAL_NORM = 0, 1, 2
Q, K, V, O, GATE, UP, DOWN, IN_NORM, POST_NORM, Q_BIAS, K_BIAS, V_BIAS = range(12)


def tensor_id(kind: int, layer: int | None = None) -> int:
    return kind if layer is None else (layer + 1) * 16 + kind


def _splitmix64(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def stream_key(seed: int, tid: int) -> np.uint64:
    with np.errstate(over="ignore"):
        z = np.array([(seed + 0x100000001B3 * tid) & 0xFFFFFFFFFFFFFFFF], dtype=np.uint64)
        return _splitmix64(z)[0]


def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even fp32 -> bf16 bit pattern (uint16). No NaN handling needed here."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    r = u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))
    return (r >> np.uint32(16)).astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (b.astype(np.uint32) << np.uint32(16)).view(np.float32)


def round_bf16(x: np.ndarray) -> np.ndarray:
    return bf16_bits_to_f32(f32_to_bf16_bits(x))


_clib = None


def _c_helper():
    """oracle/_build/liboracle_wgen.so (oracle/csrc/wgen.c) when it has been built; else None."""
    global _clib
    if _clib is None:
        import ctypes as C
        import os
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_build", "liboracle_wgen.so")
        if os.path.exists(path):
            lib = C.CDLL(path)
            lib.oracle_gen.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
            lib.oracle_gen.restype = None
            _clib = lib
        else:
            _clib = False
    return _clib or None


def gen_bits_numpy(seed: int, tid: int, start: int, count: int, std: float = 0.02, offset: float = 0.0) -> np.ndarray:
    """Pure-numpy definition of the generator (the C helper must equal it bit for bit)."""
    if count > (1 << 24):
        parts = [gen_bits_numpy(seed, tid, s, min(1 << 24, start + count - s), std, offset)
                 for s in range(start, start + count, 1 << 24)]
        return np.concatenate(parts)
    key = stream_key(seed, tid)
    with np.errstate(over="ignore"):
        idx = np.arange(start, start + count, dtype=np.uint64) + key
    h = _splitmix64(idx)
    m = np.uint64(0xFFFF)
    s = (h & m) + ((h >> np.uint64(16)) & m) + ((h >> np.uint64(32)) & m) + (h >> np.uint64(48))
    u = (s.astype(np.int64) - 131070).astype(np.float32)
    scale = np.float32(std / IH4_STD)
    v = u * scale                      # one correctly rounded fp32 multiply
    if offset != 0.0:
        v = v + np.float32(offset)     # one correctly rounded fp32 add
    return f32_to_bf16_bits(v)


def gen_bits(seed: int, tid: int, start: int, count: int, std: float = 0.02, offset: float = 0.0) -> np.ndarray:
    """bf16 bit patterns of elements [start, start+count) of tensor ``tid``."""
    lib = _c_helper()
    if lib is None:
        return gen_bits_numpy(seed, tid, start, count, std, offset)
    out = np.empty(count, dtype=np.uint16)
    lib.oracle_gen(int(stream_key(seed, tid)), start, count, np.float32(std / IH4_STD), np.float32(offset), None,
                   out.ctypes.data)
    return out


def gen_tensor(seed: int, kind: int, layer: int | None, shape, std: float = 0.02, offset: float = 0.0) -> np.ndarray:
    """fp32 array holding bf16-representable values, row-major in the HF tensor layout."""
    n = int(np.prod(shape))
    lib = _c_helper()
    if lib is None:
        return bf16_bits_to_f32(gen_bits_numpy(seed, tensor_id(kind, layer), 0, n, std, offset)).reshape(shape)
    out = np.empty(n, dtype=np.float32)
    lib.oracle_gen(int(stream_key(seed, tensor_id(kind, layer))), 0, n, np.float32(std / IH4_STD), np.float32(offset),
                   out.ctypes.data, None)
    return out.reshape(shape)


def norm_tensor(seed: int, kind: int, layer: int | None, n: int) -> np.ndarray:
    return gen_tensor(seed, kind, layer, (n,), std=0.02, offset=1.0)
