"""CPU restatement of the engine's sampler (TEST INFRASTRUCTURE; see oracle/__init__.py).

The reference sends no sampling parameter (/root/reference/src/adapters/local-llm.ts:162-165), so there is nothing in
it to pin this against ("parity unpinned"); the FILTER semantics are those of Hugging Face transformers 5.5
``TopKLogitsWarper`` followed by ``TopPLogitsWarper`` (generation/logits_process.py), against which
tests/test_oracle_sampling.py checks ``keep_mask`` on tie-free inputs.  Two deliberate differences, both stated in
theroundtaible_b200/csrc/sampler.cuh: ties at a threshold are all kept (HF's sort splits them arbitrarily), and
probability mass is summed in 2^-40 fixed point so that the result does not depend on summation order.

The draw is Gumbel-max with a counter-based noise stream keyed on (seed, step, slot, token index), restated bit for
bit from csrc/gemv.cuh ``gumbel_noise`` up to the two float32 logarithms.
"""
from __future__ import annotations

import numpy as np

_M64 = (1 << 64) - 1


def gumbel_noise(seed: int, step: int, slot: int, idx: np.ndarray) -> np.ndarray:
    """float32 Gumbel(0,1) noise for token indices ``idx`` (csrc/gemv.cuh gumbel_noise)."""
    idx = np.asarray(idx, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = (np.uint64(seed & _M64) + np.uint64(0x9E3779B97F4A7C15) * np.uint64((step & 0xFFFFFFFF) * 64 + (slot & 0xFFFFFFFF))
             + (idx << np.uint64(20)))
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    u = ((z >> np.uint64(40)).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 16777216.0)
    return -np.log(-np.log(u, dtype=np.float32), dtype=np.float32)


def keep_mask(logits: np.ndarray, temperature: float, top_k: int = 0, top_p: float = 0.0):
    """(mask [V] bool, threshold logit, margin): tokens that survive top-k then top-p.  ``margin`` is the relative
    distance between the probability mass at the top-p cut and its target -- a caller comparing against another
    implementation skips cases where it is below that implementation's rounding noise."""
    l = np.asarray(logits, dtype=np.float32)
    V = l.shape[0]
    inv_temp = np.float32(1.0) / np.float32(temperature)
    thresh = -np.inf
    margin = np.inf
    if 0 < top_k < V:
        thresh = float(np.partition(l, V - top_k)[V - top_k])            # the k-th largest; ties at it are kept
    mask = l >= thresh
    if 0.0 < top_p < 1.0:
        m = l.max()
        w = np.exp((l - m) * inv_temp, dtype=np.float32)
        q = (w * np.float32(1099511627776.0)).astype(np.uint64)          # 2^-40 fixed point, truncated
        q = np.where(mask, q, np.uint64(0))
        mass = int(q.sum(dtype=np.uint64))
        target = max(1, int(float(np.float32(top_p)) * float(mass)))
        vals = np.unique(l[mask])[::-1]                                   # distinct kept values, descending
        order = np.argsort(-l, kind="stable")
        cum = np.cumsum(q[order].astype(object))                          # exact Python integers
        lsorted = l[order]
        # the cumulative mass down to and including every element equal to v
        ends = np.searchsorted(-lsorted, -vals, side="right") - 1
        cum_at = np.array([int(cum[e]) for e in ends], dtype=object)
        j = next(i for i, c in enumerate(cum_at) if c >= target)
        thresh = float(vals[j])
        below = int(cum_at[j - 1]) if j > 0 else 0
        margin = min(int(cum_at[j]) - target, target - below) / max(mass, 1)
        mask = l >= thresh
    return mask, thresh, margin


def sample(logits: np.ndarray, temperature: float, top_k: int = 0, top_p: float = 0.0, seed: int = 0, step: int = 0,
           slot: int = 0):
    """(token, top-2 margin of the noisy scores): Gumbel-max over the kept set, lowest index winning ties."""
    l = np.asarray(logits, dtype=np.float32)
    mask, _, _ = keep_mask(l, temperature, top_k, top_p)
    idx = np.nonzero(mask)[0]
    inv_temp = np.float32(1.0) / np.float32(temperature)
    score = l[idx] * inv_temp + gumbel_noise(seed, step, slot, idx)
    best = int(np.argmax(score))                                           # first maximum = lowest index
    top2 = np.partition(score, -2)[-2:] if score.size > 1 else np.array([-np.inf, score[0]])
    return int(idx[best]), float(top2[1] - top2[0])
