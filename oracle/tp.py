"""Tensor-parallel restatement of the oracle forward over torch.distributed (TEST INFRASTRUCTURE; see
oracle/__init__.py).  The reference has no collective; this states the PARTITIONING the engine
uses so that it can be checked on CPU with the gloo backend, world_size 2:

  q/k/v, gate, up   column-parallel: rank r holds heads / ffn rows [r*n/tp, (r+1)*n/tp)
  o, down           row-parallel: rank r holds the matching input columns; partial outputs are all-reduced (sum)
  lm_head           vocab-parallel: rank r holds vocab rows [r*V/tp, (r+1)*V/tp); arg-max = max over ranks of
                    (local max, global index), lowest index winning ties
"""
from __future__ import annotations

import math

import torch
import torch.distributed as dist

from .model import KVState, OracleModel, _rbf16, _rf16


@torch.no_grad()
def forward_tp(m: OracleModel, ids, kv: KVState, rank: int, tp: int):
    """Sharded forward of OracleModel.forward(round_kv=True, round_act=True).  Returns the LOCAL logits shard
    [T, vocab/tp]; kv holds only this rank's kv heads."""
    a, sd = m.a, m.sd
    ids = torch.as_tensor(ids, dtype=torch.long)
    T, p0 = ids.numel(), kv.length
    pos = torch.arange(p0, p0 + T)
    nq, nkv, F, V, hd = a.n_q // tp, a.n_kv // tp, a.ffn // tp, a.vocab // tp, a.head_dim
    G = a.n_q // a.n_kv
    x = sd["embed"][ids]
    scale = 1.0 / math.sqrt(hd)
    if not kv.k:
        kv.k = [torch.zeros(nkv, 0, hd) for _ in range(a.layers)]
        kv.v = [torch.zeros(nkv, 0, hd) for _ in range(a.layers)]
    rows = lambda t, n: t[rank * n:(rank + 1) * n]
    for l in range(a.layers):
        p = f"l{l}."
        h, r = m._rms_split(x, sd[p + "in_norm"])
        h = _rf16(h)
        q = (h @ rows(sd[p + "q"], nq * hd).T) * r
        k = (h @ rows(sd[p + "k"], nkv * hd).T) * r
        v = (h @ rows(sd[p + "v"], nkv * hd).T) * r
        if a.qkv_bias:
            q, k, v = q + rows(sd[p + "q_bias"], nq * hd), k + rows(sd[p + "k_bias"], nkv * hd), v + rows(sd[p + "v_bias"], nkv * hd)
        q = _rbf16(m._rope(q.view(T, nq, hd), pos))
        k = _rbf16(m._rope(k.view(T, nkv, hd), pos))
        v = _rbf16(v.view(T, nkv, hd))
        kv.k[l] = torch.cat([kv.k[l], k.transpose(0, 1)], dim=1)
        kv.v[l] = torch.cat([kv.v[l], v.transpose(0, 1)], dim=1)
        K, Vv = kv.k[l], kv.v[l]
        L = K.shape[1]
        qh = q.transpose(0, 1).reshape(nkv, G * T, hd)
        s = (torch.bmm(qh, K.transpose(1, 2)) * scale).view(nkv, G, T, L)
        causal = torch.arange(L)[None, :] <= pos[:, None]
        s = s.masked_fill(~causal[None, None], float("-inf"))
        o = torch.bmm(torch.softmax(s, dim=-1).view(nkv, G * T, L), Vv).view(nkv, G, T, hd)
        o = _rf16(o.permute(2, 0, 1, 3).reshape(T, nq * hd))
        part = o @ sd[p + "o"][:, rank * nq * hd:(rank + 1) * nq * hd].T          # row-parallel partial sum
        dist.all_reduce(part)
        x = x + part
        h, r = m._rms_split(x, sd[p + "post_norm"])
        h = _rf16(h)
        act = torch.nn.functional.silu((h @ rows(sd[p + "gate"], F).T) * r) * ((h @ rows(sd[p + "up"], F).T) * r)
        part = _rf16(act) @ sd[p + "down"][:, rank * F:(rank + 1) * F].T
        dist.all_reduce(part)
        x = x + part
    kv.length = p0 + T
    xf, r = m._rms_split(x, sd["final_norm"])
    return (_rf16(xf) @ rows(sd["lm_head"], V).T) * r


def argmax_tp(local_logits: torch.Tensor, rank: int, tp: int) -> list:
    """Vocab-parallel arg-max: all-gather (max value, global index) pairs; lowest index wins ties."""
    V = local_logits.shape[-1]
    val, idx = local_logits.max(-1)
    pair = torch.stack([val.double(), (idx + rank * V).double()], dim=-1)
    allp = [torch.zeros_like(pair) for _ in range(tp)]
    dist.all_gather(allp, pair)
    out = []
    for t in range(local_logits.shape[0]):
        best = max(((float(p[t, 0]), -int(p[t, 1])) for p in allp))
        out.append(-best[1])
    return out
