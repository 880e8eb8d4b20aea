"""Pins oracle/sampling.py's filters against Hugging Face transformers' TopKLogitsWarper + TopPLogitsWarper (the
published semantics csrc/sampler.cuh follows) on tie-free logits, and checks the noise stream's basic statistics."""
import numpy as np
import pytest
import torch

from oracle import sampling as S


@pytest.mark.parametrize("V,temp,top_k,top_p", [(1000, 1.0, 50, 0.0), (1000, 0.7, 0, 0.9), (5000, 1.3, 40, 0.8),
                                                (333, 1.0, 1, 0.0), (2048, 0.5, 0, 0.3), (2048, 1.0, 300, 0.999)])
def test_keep_mask_matches_hf_warpers(V, temp, top_k, top_p):
    from transformers.generation.logits_process import TemperatureLogitsWarper, TopKLogitsWarper, TopPLogitsWarper
    rng = np.random.default_rng(V + top_k)
    checked = 0
    for _ in range(8):
        logits = (rng.standard_normal(V) * 3.0).astype(np.float32)
        while len(np.unique(logits)) != V:                   # HF splits ties by sort order; this restatement keeps them all
            logits = (rng.standard_normal(V) * 3.0).astype(np.float32)
        mask, thresh, margin = S.keep_mask(logits, temp, top_k, top_p)
        if margin < 1e-5:
            continue                                         # the cut sits within float32 softmax rounding of its target
        scores = TemperatureLogitsWarper(temp)(None, torch.from_numpy(logits)[None])
        if top_k:
            scores = TopKLogitsWarper(top_k)(None, scores)
        if top_p:
            scores = TopPLogitsWarper(top_p)(None, scores)
        want = torch.isfinite(scores[0]).numpy()
        assert np.array_equal(mask, want)
        assert thresh == logits[mask].min()
        checked += 1
    assert checked >= 6


def test_ties_at_the_threshold_are_all_kept():
    logits = np.array([3.0, 1.0, 2.0, 2.0, 2.0, 0.5], dtype=np.float32)
    mask, thresh, _ = S.keep_mask(logits, 1.0, top_k=2)
    assert thresh == 2.0 and mask.tolist() == [True, False, True, True, True, False]
    mask, thresh, _ = S.keep_mask(logits, 1.0, top_p=0.5)     # p(3.0) = 0.43 < 0.5: the cut falls on the tied 2.0s
    assert thresh == 2.0 and mask.sum() == 4
    mask, _, _ = S.keep_mask(logits, 1.0, top_p=0.4)
    assert mask.tolist() == [True, False, False, False, False, False]
    mask, _, _ = S.keep_mask(logits, 1.0)                     # no filter
    assert mask.all()


def test_gumbel_noise_stream():
    g = S.gumbel_noise(5, 3, 1, np.arange(200000))
    assert g.dtype == np.float32 and np.isfinite(g).all()
    assert abs(g.mean() - 0.5772) < 0.01 and abs(g.var() - np.pi ** 2 / 6) < 0.03      # Gumbel(0, 1) moments
    assert not np.array_equal(g[:100], S.gumbel_noise(5, 4, 1, np.arange(100)))
    assert not np.array_equal(g[:100], S.gumbel_noise(5, 3, 2, np.arange(100)))
    assert np.array_equal(g[:100], S.gumbel_noise(5, 3, 1, np.arange(100)))


def test_sample_follows_the_renormalised_distribution():
    logits = np.log(np.array([0.4, 0.3, 0.15, 0.1, 0.05], dtype=np.float32))
    counts = np.zeros(5)
    for step in range(4000):
        tok, _ = S.sample(logits, 1.0, top_k=3, seed=9, step=step)
        counts[tok] += 1
    assert counts[3] == 0 and counts[4] == 0
    want = np.array([0.4, 0.3, 0.15]) / 0.85
    assert np.abs(counts[:3] / 4000 - want).max() < 0.03
