"""ctypes binding of include/rt_engine.h (the C-ABI a Node-API addon would bind).

There is no CPU fallback: if the shared library is missing or no B200 is present, construction
raises.  Nothing here imports ``oracle``.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_lib", "librt_engine.so")

RT_FLAG_NO_GRAPH = 1
RT_FLAG_NO_PREFIX_SHARE = 2
RT_FLAG_PREFILL_VIA_DECODE = 4
RT_FLAG_PREFIX_CACHE = 8

RT_ERR_TIMEOUT = -4
RT_ERR_OOM = -3


class RtConfig(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("family", C.c_int32), ("layers", C.c_int32), ("hidden", C.c_int32),
        ("n_q", C.c_int32), ("n_kv", C.c_int32), ("head_dim", C.c_int32), ("ffn", C.c_int32), ("vocab", C.c_int32),
        ("tied", C.c_int32), ("qkv_bias", C.c_int32), ("rope_theta", C.c_float), ("rms_eps", C.c_float),
        ("rope_factor", C.c_float), ("rope_low_freq", C.c_float), ("rope_high_freq", C.c_float),
        ("rope_orig_ctx", C.c_int32), ("seed", C.c_uint64), ("weights_path", C.c_char_p), ("device", C.c_int32),
        ("tp_rank", C.c_int32), ("tp_size", C.c_int32), ("nccl_id", C.c_void_p), ("max_batch", C.c_int32),
        ("max_ctx", C.c_int32), ("kv_pool_bytes", C.c_int64), ("flags", C.c_int32), ("n_eos", C.c_int32),
        ("eos_ids", C.c_int32 * 3), ("reserved", C.c_int32),
    ]


class RtSampling(C.Structure):
    _fields_ = [("max_new_tokens", C.c_int32), ("temperature", C.c_float), ("top_k", C.c_int32), ("top_p", C.c_float),
                ("seed", C.c_uint64), ("ignore_eos", C.c_int32), ("reserved", C.c_int32)]


class RtResult(C.Structure):
    _fields_ = [("text", C.POINTER(C.c_char)), ("text_len", C.c_size_t), ("tokens", C.POINTER(C.c_int32)),
                ("n_tokens", C.c_int32), ("n_prompt_tokens", C.c_int32), ("prefix_shared_tokens", C.c_int32),
                ("finish_reason", C.c_int32), ("prefill_ms", C.c_double), ("decode_ms", C.c_double),
                ("decode_steps", C.c_int32), ("gpu_launches", C.c_int32), ("prefix_cached_tokens", C.c_int32),
                ("reserved", C.c_int32)]


class RtModelInfo(C.Structure):
    _fields_ = [("context_length", C.c_int32), ("vocab", C.c_int32), ("n_params", C.c_int64),
                ("weight_bytes_per_step", C.c_int64), ("kv_bytes_per_token", C.c_int64), ("max_batch", C.c_int32),
                ("page_size", C.c_int32), ("n_pages", C.c_int32), ("sm_count", C.c_int32)]


EXPORTS = ("rt_arch_by_name", "rt_engine_create", "rt_engine_destroy", "rt_last_error", "rt_model_info_get",
           "rt_generate", "rt_generate_batch", "rt_generate_ids", "rt_seq_reset", "rt_step_logits", "rt_result_free",
           "rt_tokenize_chat", "rt_detokenize", "rt_nccl_unique_id", "rt_weights_export", "rt_k_gemv", "rt_k_rmsnorm",
           "rt_k_attention", "rt_k_gemv_bench", "rt_k_attention_noalloc", "rt_k_gemm_tc5", "rt_k_sample_topkp")

_lib = None


class EngineError(RuntimeError):
    """Error raised by the native backend.  ``str(err)`` is what the adapter would throw, so that the
    reference's substring classifier (src/utils/errors.ts:86-126) maps it: "timed out" -> timeout,
    "not found" -> not_installed, everything else -> unknown."""

    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("RT_LIB") or _LIB_PATH        # RT_LIB: a profiling build of the same sources (build.py --trace)
    if not os.path.exists(path):
        raise EngineError(-2, f"native library not found at {path}: run `python -m theroundtaible_b200.build` "
                              "(this backend has no CPU fallback)")
    lib = C.CDLL(path)
    vp, i32, i64, u64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_size_t
    lib.rt_arch_by_name.argtypes = [C.c_char_p, C.POINTER(RtConfig)]
    lib.rt_engine_create.argtypes = [C.POINTER(RtConfig), C.POINTER(vp)]
    lib.rt_engine_destroy.argtypes = [vp]
    lib.rt_engine_destroy.restype = None
    lib.rt_last_error.argtypes = [vp]
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_model_info_get.argtypes = [vp, C.POINTER(RtModelInfo)]
    lib.rt_generate.argtypes = [vp, C.c_char_p, sz, C.POINTER(RtSampling), u64, C.POINTER(RtResult)]
    lib.rt_generate_batch.argtypes = [vp, C.POINTER(C.c_char_p), C.POINTER(sz), i32, C.POINTER(RtSampling), u64,
                                      C.POINTER(RtResult)]
    lib.rt_generate_ids.argtypes = [vp, C.POINTER(C.POINTER(i32)), C.POINTER(i32), i32, C.POINTER(RtSampling), u64,
                                    C.POINTER(RtResult)]
    lib.rt_seq_reset.argtypes = [vp]
    lib.rt_step_logits.argtypes = [vp, C.POINTER(i32), i32, i32, C.POINTER(C.c_float)]
    lib.rt_result_free.argtypes = [C.POINTER(RtResult)]
    lib.rt_result_free.restype = None
    lib.rt_tokenize_chat.argtypes = [C.c_char_p, sz, C.POINTER(i32), i32]
    lib.rt_detokenize.argtypes = [C.POINTER(i32), i32, C.c_char_p, i64]
    lib.rt_detokenize.restype = i64
    lib.rt_nccl_unique_id.argtypes = [vp]
    lib.rt_weights_export.argtypes = [vp, i32, C.POINTER(C.c_uint16), i64]
    lib.rt_k_gemv.argtypes = [C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), C.POINTER(C.c_float), i32, i32, i32, i32]
    lib.rt_k_rmsnorm.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), i32, i32,
                                 C.c_float, i32]
    lib.rt_k_attention.argtypes = [C.POINTER(C.c_uint16)] * 3 + [C.POINTER(i32), C.POINTER(C.c_float)] + [i32] * 8
    lib.rt_k_gemv_bench.argtypes = [i32, i32, i32, i32, i32, i32, C.POINTER(C.c_double)]
    lib.rt_k_gemm_tc5.argtypes = [C.POINTER(C.c_uint16), C.POINTER(C.c_uint16), C.POINTER(C.c_float), i32, i32, i32, i32]
    lib.rt_k_sample_topkp.argtypes = [C.POINTER(C.c_float), i32, i32, C.c_float, i32, C.c_float, C.c_uint64, i32,
                                      C.POINTER(i32), C.POINTER(i32), C.POINTER(C.c_float), i32]
    lib.rt_k_attention_noalloc.argtypes = [vp, vp, vp, vp, i32, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]
    _lib = lib
    return lib


def _check(rc: int, handle=None):
    if rc != 0:
        msg = load_library().rt_last_error(handle)
        raise EngineError(rc, (msg or b"unknown error").decode("utf-8", "replace"))


def arch_config(model: str) -> RtConfig:
    cfg = RtConfig()
    _check(load_library().rt_arch_by_name(model.encode(), C.byref(cfg)))
    return cfg


def config_from_hf(hf) -> RtConfig:
    """Architecture half of rt_config from a Hugging Face ``config.json`` (a dict, a PretrainedConfig, or the path of
    a checkpoint directory / config file).  Llama and Qwen2 only; anything this backend has no kernel
    for (o_proj / MLP biases, sliding windows, other rope types) is refused here rather than silently ignored."""
    import json
    import os
    if isinstance(hf, (str, os.PathLike)):
        path = os.fspath(hf)
        if os.path.isdir(path):
            path = os.path.join(path, "config.json")
        with open(path) as f:
            hf = json.load(f)
    elif hasattr(hf, "to_dict"):
        hf = hf.to_dict()
    mt = hf.get("model_type")
    if mt not in ("llama", "qwen2"):
        raise EngineError(-1, f"model_type '{mt}' not found among the supported families (llama, qwen2)")
    if mt == "llama" and (hf.get("attention_bias") or hf.get("mlp_bias")):
        raise EngineError(-1, "llama checkpoints with attention_bias / mlp_bias are not supported")
    if mt == "qwen2" and hf.get("use_sliding_window"):
        raise EngineError(-1, "qwen2 sliding-window attention is not supported")
    cfg = RtConfig()
    cfg.struct_size = C.sizeof(RtConfig)
    cfg.family = 0 if mt == "llama" else 1
    cfg.layers, cfg.hidden = hf["num_hidden_layers"], hf["hidden_size"]
    cfg.n_q = hf["num_attention_heads"]
    cfg.n_kv = hf.get("num_key_value_heads") or cfg.n_q
    cfg.head_dim = hf.get("head_dim") or cfg.hidden // cfg.n_q
    cfg.ffn, cfg.vocab = hf["intermediate_size"], hf["vocab_size"]
    cfg.tied = 1 if hf.get("tie_word_embeddings") else 0
    cfg.qkv_bias = 1 if mt == "qwen2" else 0
    cfg.rms_eps = hf.get("rms_norm_eps", 1e-6)
    rope = hf.get("rope_parameters") or hf.get("rope_scaling") or {}
    cfg.rope_theta = rope.get("rope_theta") or hf.get("rope_theta") or 10000.0
    rtype = rope.get("rope_type") or rope.get("type") or "default"
    if rtype == "llama3":
        cfg.rope_factor = rope["factor"]
        cfg.rope_low_freq, cfg.rope_high_freq = rope["low_freq_factor"], rope["high_freq_factor"]
        cfg.rope_orig_ctx = rope["original_max_position_embeddings"]
    elif rtype != "default":
        raise EngineError(-1, f"rope type '{rtype}' is not supported (default and llama3 are)")
    eos = hf.get("eos_token_id")
    eos = [] if eos is None else ([eos] if isinstance(eos, int) else list(eos))
    cfg.n_eos = min(len(eos), 3)
    for i in range(cfg.n_eos):
        cfg.eos_ids[i] = int(eos[i])
    return cfg


@dataclass
class Generation:
    text: str
    tokens: list
    n_prompt_tokens: int
    prefix_shared_tokens: int
    finish_reason: int
    prefill_ms: float
    decode_ms: float
    decode_steps: int
    gpu_launches: int
    prefix_cached_tokens: int = 0


def _p(arr: np.ndarray, ctype):
    return arr.ctypes.data_as(C.POINTER(ctype))


class Engine:
    """One engine per GPU.  Calls are serialised by the caller (the reference issues one execute() at a time)."""

    def __init__(self, model, *, seed: int = 0, device: int = 0, max_batch: int = 8, max_ctx: int = 4096,
                 kv_pool_bytes: int = 0, flags: int = 0, tp_rank: int = 0, tp_size: int = 1, nccl_id: bytes | None = None,
                 weights_path: str | None = None):
        """``model`` is a name rt_arch_by_name knows (seeded random-init weights unless ``weights_path`` is given) or an
        RtConfig (e.g. from ``config_from_hf``).  ``weights_path``: a .safetensors file or a directory of them."""
        lib = load_library()
        if isinstance(model, RtConfig):
            cfg, model = model, "custom"
        else:
            cfg = arch_config(model)
        cfg.seed = seed
        self._weights_path = os.fspath(weights_path).encode() if weights_path else None    # must outlive rt_engine_create
        cfg.weights_path = self._weights_path
        cfg.device = device
        self._nccl_id = C.create_string_buffer(nccl_id, 128) if nccl_id else None    # must outlive rt_engine_create
        cfg.tp_rank, cfg.tp_size = tp_rank, tp_size
        cfg.nccl_id = C.cast(self._nccl_id, C.c_void_p) if self._nccl_id else None
        cfg.max_batch, cfg.max_ctx, cfg.kv_pool_bytes, cfg.flags = max_batch, max_ctx, kv_pool_bytes, flags
        self.cfg = cfg
        self.model = model
        self._h = C.c_void_p()
        _check(lib.rt_engine_create(C.byref(cfg), C.byref(self._h)))
        self._lib = lib
        info = RtModelInfo()
        _check(lib.rt_model_info_get(self._h, C.byref(info)), self._h)
        self.info = info
        self.vocab = info.vocab

    @classmethod
    def from_pretrained(cls, path: str, **kw) -> "Engine":
        """A Hugging Face checkpoint directory: ``config.json`` + ``*.safetensors`` (Llama / Qwen2)."""
        return cls(config_from_hf(path), weights_path=path, **kw)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.rt_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- generation ---------------------------------------------------------------------------
    @staticmethod
    def _sampling(max_new_tokens, temperature, seed, ignore_eos, top_k=0, top_p=0.0):
        return RtSampling(max_new_tokens, temperature, top_k, top_p, seed, 1 if ignore_eos else 0, 0)

    def _collect(self, res, n):
        out = []
        for i in range(n):
            r = res[i]
            toks = [r.tokens[k] for k in range(r.n_tokens)]
            text = C.string_at(r.text, r.text_len).decode("utf-8", "replace")
            out.append(Generation(text, toks, r.n_prompt_tokens, r.prefix_shared_tokens, r.finish_reason, r.prefill_ms,
                                  r.decode_ms, r.decode_steps, r.gpu_launches, r.prefix_cached_tokens))
            self._lib.rt_result_free(C.byref(res[i]))
        return out

    def generate_ids(self, prompts, max_new_tokens=64, temperature=0.0, seed=0, ignore_eos=True, deadline_ms=0,
                     top_k=0, top_p=0.0):
        n = len(prompts)
        arrs = [np.ascontiguousarray(p, dtype=np.int32) for p in prompts]
        ptrs = (C.POINTER(C.c_int32) * n)(*[_p(a, C.c_int32) for a in arrs])
        lens = (C.c_int32 * n)(*[len(a) for a in arrs])
        res = (RtResult * n)()
        s = self._sampling(max_new_tokens, temperature, seed, ignore_eos, top_k, top_p)
        _check(self._lib.rt_generate_ids(self._h, ptrs, lens, n, C.byref(s), deadline_ms, res), self._h)
        return self._collect(res, n)

    def generate_batch(self, prompts, max_new_tokens=0, temperature=0.0, seed=0, ignore_eos=False, deadline_ms=0,
                       top_k=0, top_p=0.0):
        n = len(prompts)
        raw = [p.encode("utf-8") for p in prompts]
        ptrs = (C.c_char_p * n)(*raw)
        lens = (C.c_size_t * n)(*[len(b) for b in raw])
        res = (RtResult * n)()
        s = self._sampling(max_new_tokens, temperature, seed, ignore_eos, top_k, top_p)
        _check(self._lib.rt_generate_batch(self._h, ptrs, lens, n, C.byref(s), deadline_ms, res), self._h)
        return self._collect(res, n)

    def generate(self, prompt: str, **kw) -> Generation:
        return self.generate_batch([prompt], **kw)[0]

    # ---- parity hooks -------------------------------------------------------------------------
    def seq_reset(self):
        _check(self._lib.rt_seq_reset(self._h), self._h)

    def step_logits(self, ids, mode: int = 0) -> np.ndarray:
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        out = np.empty((len(ids), self.vocab // self.cfg.tp_size), dtype=np.float32)    # vocab-parallel shard under TP
        _check(self._lib.rt_step_logits(self._h, _p(ids, C.c_int32), len(ids), mode, _p(out, C.c_float)), self._h)
        return out

    def export_weight(self, tensor_id: int, shape) -> np.ndarray:
        out = np.empty(int(np.prod(shape)), dtype=np.uint16)
        _check(self._lib.rt_weights_export(self._h, tensor_id, _p(out, C.c_uint16), out.size), self._h)
        return out.reshape(shape)


# ---- tokenizer + kernel hooks (module level: they need no engine) ----------------------------------
def nccl_unique_id() -> bytes:
    """rank 0 of a tensor-parallel group calls this and ships the 128 bytes to every rank (rt_config.nccl_id)."""
    buf = C.create_string_buffer(128)
    _check(load_library().rt_nccl_unique_id(buf))
    return buf.raw


def tokenize_chat(text: str) -> list:
    lib = load_library()
    b = text.encode("utf-8")
    n = lib.rt_tokenize_chat(b, len(b), None, 0)
    arr = (C.c_int32 * n)()
    lib.rt_tokenize_chat(b, len(b), arr, n)
    return list(arr)


def detokenize(ids) -> str:
    lib = load_library()
    a = np.ascontiguousarray(ids, dtype=np.int32)
    n = lib.rt_detokenize(_p(a, C.c_int32), len(a), None, 0)
    buf = C.create_string_buffer(n + 1)
    lib.rt_detokenize(_p(a, C.c_int32), len(a), buf, n + 1)
    return buf.raw[:n].decode("utf-8", "replace")


def k_gemv(w_bits: np.ndarray, x_bits: np.ndarray, device: int = 0) -> np.ndarray:
    N, K = w_bits.shape
    T = x_bits.shape[0]
    w = np.ascontiguousarray(w_bits, dtype=np.uint16)
    x = np.ascontiguousarray(x_bits, dtype=np.uint16)
    y = np.empty((T, N), dtype=np.float32)
    _check(load_library().rt_k_gemv(_p(w, C.c_uint16), _p(x, C.c_uint16), _p(y, C.c_float), N, K, T, device))
    return y


def k_rmsnorm(x: np.ndarray, w_bits: np.ndarray, eps: float, device: int = 0) -> np.ndarray:
    T, H = x.shape
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w_bits, dtype=np.uint16)
    y = np.empty((T, H), dtype=np.uint16)
    _check(load_library().rt_k_rmsnorm(_p(x, C.c_float), _p(w, C.c_uint16), _p(y, C.c_uint16), T, H, eps, device))
    return y


def k_attention(q_bits, k_bits, v_bits, lens, shared: int = 0, chunk: int = 64, device: int = 0) -> np.ndarray:
    T, n_q, hd = q_bits.shape
    _, ctx, n_kv, _ = k_bits.shape
    q = np.ascontiguousarray(q_bits, dtype=np.uint16)
    k = np.ascontiguousarray(k_bits, dtype=np.uint16)
    v = np.ascontiguousarray(v_bits, dtype=np.uint16)
    ln = np.ascontiguousarray(lens, dtype=np.int32)
    out = np.empty((T, n_q, hd), dtype=np.float32)
    _check(load_library().rt_k_attention(_p(q, C.c_uint16), _p(k, C.c_uint16), _p(v, C.c_uint16), _p(ln, C.c_int32),
                                         _p(out, C.c_float), T, ctx, n_q, n_kv, hd, shared, chunk, device))
    return out


def k_sample_topkp(logits: np.ndarray, temperature: float, top_k: int = 0, top_p: float = 0.0, seed: int = 0, step: int = 0,
                   device: int = 0):
    """(tokens [T], kept-set sizes [T], thresholds [T]) of csrc/sampler.cuh on host logits [T][V]."""
    logits = np.ascontiguousarray(logits, dtype=np.float32)
    T, V = logits.shape
    tok, kept, th = np.zeros(T, np.int32), np.zeros(T, np.int32), np.zeros(T, np.float32)
    _check(load_library().rt_k_sample_topkp(_p(logits, C.c_float), T, V, temperature, top_k, top_p, seed, step,
                                            _p(tok, C.c_int32), _p(kept, C.c_int32), _p(th, C.c_float), device))
    return tok, kept, th


def k_gemm_tc5(a_bits: np.ndarray, w_bits: np.ndarray, device: int = 0) -> np.ndarray:
    S, K = a_bits.shape
    N = w_bits.shape[0]
    a = np.ascontiguousarray(a_bits, dtype=np.uint16)
    w = np.ascontiguousarray(w_bits, dtype=np.uint16)
    c = np.empty((S, N), dtype=np.float32)
    _check(load_library().rt_k_gemm_tc5(_p(a, C.c_uint16), _p(w, C.c_uint16), _p(c, C.c_float), S, N, K, device))
    return c


def k_gemv_bench(N: int, K: int, T: int, iters: int = 50, device: int = 0, epi: int = 0) -> float:
    """Mean ms per launch of the decode projection kernel on a matrix re-streamed from HBM; epi 2 = SwiGLU epilogue."""
    ms = C.c_double()
    _check(load_library().rt_k_gemv_bench(N, K, T, iters, epi, device, C.byref(ms)))
    return ms.value


def k_attention_noalloc(q_ptr: int, kpool_ptr: int, vpool_ptr: int, page_table_ptr: int, pt_stride: int, query_pos_ptr: int,
                        out_ptr: int, T: int, max_len: int, n_q: int, n_kv: int, hd: int, chunk: int, device: int = 0,
                        stream: int = 0) -> None:
    """rt_k_attention_noalloc on raw DEVICE addresses (ints).  For hosts that must not allocate: an embedding
    framework that owns all of HBM through its own pool, or a token loop that cannot take cudaMalloc's
    device-wide synchronisation.  Asynchronous on ``stream``; raises EngineError when the built-in 4 MiB
    split-KV scratch would not suffice."""
    _check(load_library().rt_k_attention_noalloc(q_ptr, kpool_ptr, vpool_ptr, page_table_ptr, pt_stride, query_pos_ptr,
                                                 out_ptr, T, max_len, n_q, n_kv, hd, chunk, device, stream or None))
