"""Architecture table for the oracle (TEST INFRASTRUCTURE; see oracle/__init__.py).

Public architecture constants of the model families BASELINE.json names
(/root/repo/SURVEY.md section 8a table).  The reference adapter only carries a model *name*
(/root/reference/src/types.ts:29-34 ``AdapterLocalConfig.model``); the dimensions live in
the external server it calls.
"""
from __future__ import annotations

from dataclasses import dataclass, asdict, replace


@dataclass(frozen=True)
class Arch:
    name: str
    family: str            # "llama" | "qwen2"
    layers: int
    hidden: int
    n_q: int
    n_kv: int
    head_dim: int
    ffn: int
    vocab: int
    tied: bool
    rope_theta: float
    rms_eps: float
    qkv_bias: bool
    # llama3 rope scaling (factor 0 => plain rope)
    rope_factor: float = 0.0
    rope_low_freq: float = 1.0
    rope_high_freq: float = 4.0
    rope_orig_ctx: int = 8192

    def asdict(self):
        return asdict(self)

    def with_layers(self, n):
        return replace(self, layers=n, name=f"{self.name}-L{n}")


ARCHS = {
    "qwen2.5-0.5b": Arch("qwen2.5-0.5b", "qwen2", 24, 896, 14, 2, 64, 4864, 151936, True, 1e6, 1e-6, True),
    "llama-3.2-1b": Arch("llama-3.2-1b", "llama", 16, 2048, 32, 8, 64, 8192, 128256, True, 5e5, 1e-5, False,
                         32.0, 1.0, 4.0, 8192),
    "llama-3.1-8b": Arch("llama-3.1-8b", "llama", 32, 4096, 32, 8, 128, 14336, 128256, False, 5e5, 1e-5, False,
                         8.0, 1.0, 4.0, 8192),
    "qwen2.5-32b": Arch("qwen2.5-32b", "qwen2", 64, 5120, 40, 8, 128, 27648, 152064, False, 1e6, 1e-5, True),
    "llama-3.1-70b": Arch("llama-3.1-70b", "llama", 80, 8192, 64, 8, 128, 28672, 128256, False, 5e5, 1e-5, False,
                          8.0, 1.0, 4.0, 8192),
    # tiny shapes for fast parity tests (cover: llama3 rope scaling, untied head, odd GQA ratios, biases)
    "tiny-llama": Arch("tiny-llama", "llama", 2, 128, 4, 2, 32, 256, 512, True, 5e5, 1e-5, False,
                       32.0, 1.0, 4.0, 64),
    "tiny-llama-hd128": Arch("tiny-llama-hd128", "llama", 2, 256, 4, 1, 128, 512, 640, False, 5e5, 1e-5, False,
                             8.0, 1.0, 4.0, 64),
    "tiny-qwen": Arch("tiny-qwen", "qwen2", 3, 448, 14, 2, 32, 384, 768, True, 1e6, 1e-6, True),
    "tiny-qwen-hd128": Arch("tiny-qwen-hd128", "qwen2", 2, 640, 5, 1, 128, 768, 1024, False, 1e6, 1e-5, True),
}


def param_count(a: Arch) -> int:
    per_layer = (a.hidden * (a.n_q + 2 * a.n_kv) * a.head_dim + a.n_q * a.head_dim * a.hidden
                 + 3 * a.hidden * a.ffn + 2 * a.hidden)
    if a.qkv_bias:
        per_layer += (a.n_q + 2 * a.n_kv) * a.head_dim
    n = a.layers * per_layer + a.vocab * a.hidden + a.hidden
    if not a.tied:
        n += a.vocab * a.hidden
    return n


def decode_weight_bytes(a: Arch) -> int:
    """P_dec of SURVEY.md 8(a): bf16 weight bytes one decode step must read (embedding rows excluded
    unless the table doubles as the tied lm_head)."""
    per_layer = (a.hidden * (a.n_q + 2 * a.n_kv) * a.head_dim + a.n_q * a.head_dim * a.hidden
                 + 3 * a.hidden * a.ffn + 2 * a.hidden)
    return 2 * (a.layers * per_layer + a.vocab * a.hidden + a.hidden)


def kv_bytes_per_token(a: Arch) -> int:
    return 2 * a.layers * a.n_kv * a.head_dim * 2
