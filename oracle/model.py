"""CPU restatement of the decoder-only transformer forward (TEST INFRASTRUCTURE; see
oracle/__init__.py).

This is the arithmetic that happens on the far side of the reference's ``fetch``
(/root/reference/src/adapters/local-llm.ts:159-193: POST /v1/chat/completions, read
``choices[0].message.content``).  The algorithm lives in an un-vendored, un-pinned external
server (Ollama / LM Studio), so this file restates the published Llama-3 / Qwen2 algorithm
as implemented by Hugging Face transformers 5.5.0 (models/llama/modeling_llama.py,
models/qwen2/modeling_qwen2.py, modeling_rope_utils.py:550-625) and is pinned against it by
tests/test_oracle_golden.py.

Rounding points (they mirror the engine so that parity can be tight):
  * weights are bf16 values (held as fp32 here);
  * K (after RoPE) and V are rounded to bf16 when they enter the KV cache, and every
    attention read -- including the current token's own K/V -- sees the rounded values
    (``round_kv``);
  * with ``round_act`` the inputs of every projection (x * norm weight -- the RMSNorm scalar r is applied
    to the projection's output, see ``_rms_split`` --, attention output, SwiGLU output) are rounded to
    **fp16** (11 significant bits, saturating at +-65504) and Q is rounded to bf16.  fp16 and not bf16: with
    bf16 matmul inputs two correct implementations that only differ in fp32 summation order drift
    1.2e-2 of max |logit| apart on the 16-layer Llama-3.2-1B (each rounding point turns a 1e-6
    difference into 2^-9 flips, and 64 rounding points decorrelate them), which is above the 1e-2
    parity bar; with fp16 inputs the same experiment gives 2.5e-3 (scripts/noise_floor_probe.py);
  * everything else (residual stream, softmax, accumulation) is fp32.
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass, field

import numpy as np
import torch

from . import weights as W
from .archs import Arch


def rope_inv_freq(a: Arch) -> np.ndarray:
    """fp32 inverse frequencies; computed in float64 then rounded once so that C++ (engine
    host code) and numpy agree bit for bit.  Follows modeling_rope_utils.py:606-624."""
    dim = a.head_dim
    i = np.arange(0, dim, 2, dtype=np.float64)
    inv = 1.0 / (float(a.rope_theta) ** (i / dim))
    if a.rope_factor > 0.0:
        low_wl = a.rope_orig_ctx / a.rope_low_freq
        high_wl = a.rope_orig_ctx / a.rope_high_freq
        wl = 2.0 * math.pi / inv
        inv_l = np.where(wl > low_wl, inv / a.rope_factor, inv)
        smooth = (a.rope_orig_ctx / wl - a.rope_low_freq) / (a.rope_high_freq - a.rope_low_freq)
        smoothed = (1.0 - smooth) * inv_l / a.rope_factor + smooth * inv_l
        medium = (~(wl < high_wl)) & (~(wl > low_wl))
        inv = np.where(medium, smoothed, inv_l)
    return inv.astype(np.float32)


def rope_table(a: Arch, n_pos: int):
    """cos/sin [n_pos, head_dim/2] fp32: angle = fp32(pos) * inv_freq in fp32, cos/sin in
    float64 of that fp32 angle, rounded to fp32."""
    inv = rope_inv_freq(a)
    pos = np.arange(n_pos, dtype=np.float32)
    ang = (pos[:, None] * inv[None, :]).astype(np.float32)
    return np.cos(ang.astype(np.float64)).astype(np.float32), np.sin(ang.astype(np.float64)).astype(np.float32)


def _rbf16(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


def _rf16(t: torch.Tensor) -> torch.Tensor:
    """fp16 rounding of a projection input, saturating (the engine's split_f16, csrc/common.cuh)."""
    return t.clamp(-65504.0, 65504.0).to(torch.float16).to(torch.float32)


@dataclass
class KVState:
    """Per-sequence KV cache: lists (one per layer) of [n_kv, len, head_dim] fp32 tensors."""
    k: list = field(default_factory=list)
    v: list = field(default_factory=list)
    length: int = 0

    def clone(self) -> "KVState":
        return KVState([t.clone() for t in self.k], [t.clone() for t in self.v], self.length)


class OracleModel:
    def __init__(self, arch: Arch, seed: int = 0, max_pos: int = 4096, state_dict: dict | None = None):
        self.a = a = arch
        self.seed = seed
        if state_dict is None:
            state_dict = make_state_dict(a, seed)
        self.sd = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in state_dict.items()}
        cos, sin = rope_table(a, max_pos)
        self.cos, self.sin = torch.from_numpy(cos), torch.from_numpy(sin)
        self.max_pos = max_pos
        # bench.py's CPU arm: seconds spent inside the decoder-layer loops (the part that scales with depth), so that a
        # truncated-depth run of a model too big for host RAM can be extrapolated to the full layer count
        self.layer_seconds = 0.0

    # -- building blocks ------------------------------------------------------------------
    def _rms(self, x, w):
        var = (x * x).mean(-1, keepdim=True)
        return x * torch.rsqrt(var + self.a.rms_eps) * w

    def _rms_split(self, x, w):
        """RMSNorm(x)·w factored as (x*w, r) with r = rsqrt(mean x^2 + eps): the per-row scalar r commutes
        with the projection that follows, so ``(x*w) @ W.T * r`` is the same function as
        ``_rms(x, w) @ W.T`` (modeling_llama.py LlamaRMSNorm + the Linear after it).  The engine rounds
        x*w to fp16 and applies r to the fp32 accumulators; with ``round_act`` the oracle rounds at the
        same point."""
        var = (x * x).mean(-1, keepdim=True)
        return x * w, torch.rsqrt(var + self.a.rms_eps)

    def _rope(self, t, pos):  # t [T, heads, hd]; pos LongTensor [T]
        hd = self.a.head_dim
        c = self.cos[pos][:, None, :]
        s = self.sin[pos][:, None, :]
        x1, x2 = t[..., : hd // 2], t[..., hd // 2:]
        return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1)

    # -- forward ---------------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, ids, kv: KVState, round_kv: bool = True, round_act: bool = False,
                last_only: bool = False) -> torch.Tensor:
        """Append ``ids`` to the sequence held in ``kv``; return logits [T, vocab] fp32
        (or [1, vocab] with last_only)."""
        a, sd = self.a, self.sd
        ids = torch.as_tensor(ids, dtype=torch.long)
        T = ids.numel()
        p0 = kv.length
        assert p0 + T <= self.max_pos
        pos = torch.arange(p0, p0 + T)
        x = sd["embed"][ids]
        G = a.n_q // a.n_kv
        scale = 1.0 / math.sqrt(a.head_dim)
        ra = _rf16 if round_act else (lambda t: t)
        rq = _rbf16 if round_act else (lambda t: t)
        if not kv.k:
            kv.k = [torch.zeros(a.n_kv, 0, a.head_dim) for _ in range(a.layers)]
            kv.v = [torch.zeros(a.n_kv, 0, a.head_dim) for _ in range(a.layers)]
        t_layers = time.perf_counter()
        for l in range(a.layers):
            p = f"l{l}."
            h, r = self._rms_split(x, sd[p + "in_norm"])
            h = ra(h)
            q = (h @ sd[p + "q"].T) * r
            k = (h @ sd[p + "k"].T) * r
            v = (h @ sd[p + "v"].T) * r
            if a.qkv_bias:
                q = q + sd[p + "q_bias"]
                k = k + sd[p + "k_bias"]
                v = v + sd[p + "v_bias"]
            q = self._rope(q.view(T, a.n_q, a.head_dim), pos)
            k = self._rope(k.view(T, a.n_kv, a.head_dim), pos)
            v = v.view(T, a.n_kv, a.head_dim)
            if round_kv:
                k, v = _rbf16(k), _rbf16(v)
            q = rq(q)
            kv.k[l] = torch.cat([kv.k[l], k.transpose(0, 1)], dim=1)
            kv.v[l] = torch.cat([kv.v[l], v.transpose(0, 1)], dim=1)
            K, V = kv.k[l], kv.v[l]                      # [n_kv, L, hd]
            L = K.shape[1]
            qh = q.transpose(0, 1).reshape(a.n_kv, G * T, a.head_dim)   # rows: g-major then t
            s = torch.bmm(qh, K.transpose(1, 2)) * scale                  # [n_kv, G*T, L]
            s = s.view(a.n_kv, G, T, L)
            causal = torch.arange(L)[None, :] <= pos[:, None]             # [T, L]
            s = s.masked_fill(~causal[None, None], float("-inf"))
            pr = torch.softmax(s, dim=-1).view(a.n_kv, G * T, L)
            o = torch.bmm(pr, V).view(a.n_kv, G, T, a.head_dim)
            o = o.permute(2, 0, 1, 3).reshape(T, a.n_q * a.head_dim)
            x = x + ra(o) @ sd[p + "o"].T
            h, r = self._rms_split(x, sd[p + "post_norm"])
            h = ra(h)
            g = (h @ sd[p + "gate"].T) * r
            u = (h @ sd[p + "up"].T) * r
            act = torch.nn.functional.silu(g) * u
            x = x + ra(act) @ sd[p + "down"].T
        self.layer_seconds += time.perf_counter() - t_layers
        kv.length = p0 + T
        if last_only:
            x = x[-1:]
        xf, r = self._rms_split(x, sd["final_norm"])
        return (ra(xf) @ sd["lm_head"].T) * r

    @torch.no_grad()
    def generate(self, prompt_ids, max_new_tokens: int, kv: KVState | None = None, round_act: bool = True,
                 prefill_chunk: int = 512, return_logits: bool = False):
        """Greedy decoding, EOS ignored.  ``round_act=True`` is the engine's numerics
        (fp16 projection inputs, bf16 Q / K / V).  Returns ids, or (ids, per-step logits)."""
        kv = kv if kv is not None else KVState()
        ids = list(prompt_ids)
        logits = None
        for s in range(0, len(ids), prefill_chunk):
            logits = self.forward(ids[s:s + prefill_chunk], kv, round_act=round_act, last_only=True)
        out, all_logits = [], []
        for i in range(max_new_tokens):
            if return_logits:
                all_logits.append(logits[-1].clone())
            nxt = int(torch.argmax(logits[-1]))
            out.append(nxt)
            if i + 1 < max_new_tokens:
                logits = self.forward([nxt], kv, round_act=round_act, last_only=True)
        return (out, all_logits) if return_logits else out


def make_state_dict(a: Arch, seed: int) -> dict:
    """All tensors in HF layout ([out_features, in_features]); fp32 arrays of bf16 values."""
    sd = {}
    H, hd = a.hidden, a.head_dim
    sd["embed"] = W.gen_tensor(seed, W.EMBED, None, (a.vocab, H))
    sd["lm_head"] = sd["embed"] if a.tied else W.gen_tensor(seed, W.LM_HEAD, None, (a.vocab, H))
    sd["final_norm"] = W.norm_tensor(seed, W.FINAL_NORM, None, H)
    for l in range(a.layers):
        p = f"l{l}."
        sd[p + "q"] = W.gen_tensor(seed, W.Q, l, (a.n_q * hd, H))
        sd[p + "k"] = W.gen_tensor(seed, W.K, l, (a.n_kv * hd, H))
        sd[p + "v"] = W.gen_tensor(seed, W.V, l, (a.n_kv * hd, H))
        sd[p + "o"] = W.gen_tensor(seed, W.O, l, (H, a.n_q * hd))
        sd[p + "gate"] = W.gen_tensor(seed, W.GATE, l, (a.ffn, H))
        sd[p + "up"] = W.gen_tensor(seed, W.UP, l, (a.ffn, H))
        sd[p + "down"] = W.gen_tensor(seed, W.DOWN, l, (H, a.ffn))
        sd[p + "in_norm"] = W.norm_tensor(seed, W.IN_NORM, l, H)
        sd[p + "post_norm"] = W.norm_tensor(seed, W.POST_NORM, l, H)
        if a.qkv_bias:
            sd[p + "q_bias"] = W.gen_tensor(seed, W.Q_BIAS, l, (a.n_q * hd,))
            sd[p + "k_bias"] = W.gen_tensor(seed, W.K_BIAS, l, (a.n_kv * hd,))
            sd[p + "v_bias"] = W.gen_tensor(seed, W.V_BIAS, l, (a.n_kv * hd,))
    return sd


@torch.no_grad()
def decode_step_batch(m: OracleModel, tokens, kvs, round_act: bool = True) -> torch.Tensor:
    """One decode step for B independent sequences with the projections batched over B (what a CPU
    server does for concurrent requests); attention stays per sequence.  Returns logits [B, vocab].
    Numerically the same as B calls of ``forward([tok], kv)`` up to fp32 summation order."""
    a, sd = m.a, m.sd
    B = len(tokens)
    ra = _rf16 if round_act else (lambda t: t)
    rq = _rbf16 if round_act else (lambda t: t)
    pos = torch.tensor([kv.length for kv in kvs])
    x = sd["embed"][torch.as_tensor(tokens, dtype=torch.long)]
    G = a.n_q // a.n_kv
    scale = 1.0 / math.sqrt(a.head_dim)
    t_layers = time.perf_counter()
    for l in range(a.layers):
        p = f"l{l}."
        h, r = m._rms_split(x, sd[p + "in_norm"])
        h = ra(h)
        q, k, v = (h @ sd[p + "q"].T) * r, (h @ sd[p + "k"].T) * r, (h @ sd[p + "v"].T) * r
        if a.qkv_bias:
            q, k, v = q + sd[p + "q_bias"], k + sd[p + "k_bias"], v + sd[p + "v_bias"]
        q = rq(m._rope(q.view(B, a.n_q, a.head_dim), pos))
        k = _rbf16(m._rope(k.view(B, a.n_kv, a.head_dim), pos))
        v = _rbf16(v.view(B, a.n_kv, a.head_dim))
        outs = []
        for b, kv in enumerate(kvs):
            kv.k[l] = torch.cat([kv.k[l], k[b][:, None, :]], dim=1)
            kv.v[l] = torch.cat([kv.v[l], v[b][:, None, :]], dim=1)
            qh = q[b].view(a.n_kv, G, a.head_dim)
            s = torch.bmm(qh, kv.k[l].transpose(1, 2)) * scale
            o = torch.bmm(torch.softmax(s, dim=-1), kv.v[l])
            outs.append(o.reshape(a.n_q * a.head_dim))
        x = x + ra(torch.stack(outs)) @ sd[p + "o"].T
        h, r = m._rms_split(x, sd[p + "post_norm"])
        h = ra(h)
        act = torch.nn.functional.silu((h @ sd[p + "gate"].T) * r) * ((h @ sd[p + "up"].T) * r)
        x = x + ra(act) @ sd[p + "down"].T
    m.layer_seconds += time.perf_counter() - t_layers
    for kv in kvs:
        kv.length += 1
    xf, r = m._rms_split(x, sd["final_norm"])
    return (ra(xf) @ sd["lm_head"].T) * r
