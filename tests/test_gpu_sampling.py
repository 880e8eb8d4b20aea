"""top-k / top-p sampling on the GPU (csrc/sampler.cuh through rt_k_sample_topkp and through rt_generate_ids) against
oracle/sampling.py."""
import numpy as np
import pytest

from oracle import sampling as S
from theroundtaible_b200 import engine as E

pytestmark = pytest.mark.gpu


def _logits(T, V, seed, scale=3.0):
    return (np.random.default_rng(seed).standard_normal((T, V)) * scale).astype(np.float32)


@pytest.mark.parametrize("V,temp,top_k,top_p", [(128256, 1.0, 0, 0.0), (128256, 0.8, 50, 0.0), (151936, 1.0, 0, 0.9),
                                                (128256, 0.7, 40, 0.8), (4099, 1.5, 1, 0.0), (4099, 1.0, 4098, 0.0),
                                                (32000, 0.3, 0, 0.5), (32000, 1.0, 1000, 0.9999), (517, 1.0, 0, 1e-6)])
def test_kernel_matches_oracle(V, temp, top_k, top_p):
    T = 5
    logits = _logits(T, V, V + top_k)
    tok, kept, th = E.k_sample_topkp(logits, temp, top_k, top_p, seed=77, step=3)
    for t in range(T):
        mask, thresh, margin = S.keep_mask(logits[t], temp, top_k, top_p)
        if margin >= 1e-5:
            assert kept[t] == mask.sum(), (t, kept[t], mask.sum())
            assert th[t] == np.float32(thresh) or (not np.isfinite(thresh) and not np.isfinite(th[t]))
        assert mask[tok[t]] or margin < 1e-5
        want, gap = S.sample(logits[t], temp, top_k, top_p, seed=77, step=3, slot=t)
        if gap > 1e-4 and margin >= 1e-5:
            assert tok[t] == want
    if top_k == 1 or top_p == 1e-6:
        assert np.array_equal(tok, logits.argmax(-1))


def test_ties_and_degenerate_rows():
    V = 1000
    logits = np.round(_logits(4, V, 1) * 2) / 2                       # many exact ties
    logits[3] = 0.0                                                    # a constant row: everything ties
    tok, kept, th = E.k_sample_topkp(logits, 1.0, top_k=10, seed=1)
    for t in range(4):
        mask, thresh, _ = S.keep_mask(logits[t], 1.0, top_k=10)
        assert kept[t] == mask.sum() and kept[t] >= 10 and th[t] == np.float32(thresh) and mask[tok[t]]
    assert kept[3] == V
    neg = -np.abs(_logits(2, V, 2)) - 5.0                              # all-negative logits (sign handling of the radix keys)
    tok, kept, th = E.k_sample_topkp(neg, 1.0, top_k=7, top_p=0.6, seed=2)
    for t in range(2):
        mask, thresh, margin = S.keep_mask(neg[t], 1.0, 7, 0.6)
        assert margin < 1e-5 or (kept[t] == mask.sum() and th[t] == np.float32(thresh))


def test_draws_follow_the_renormalised_distribution_and_are_seeded():
    p = np.array([0.35, 0.25, 0.2, 0.1, 0.06, 0.04], dtype=np.float32)
    logits = np.tile(np.log(p), (32, 1))
    counts = np.zeros(6)
    for step in range(150):
        tok, kept, _ = E.k_sample_topkp(logits, 1.0, top_p=0.75, seed=4, step=step)     # 0.35+0.25+0.2 = 0.8 >= 0.75
        assert (kept == 3).all()
        counts += np.bincount(tok, minlength=6)
    assert counts[3:].sum() == 0
    assert np.abs(counts[:3] / counts.sum() - p[:3] / p[:3].sum()).max() < 0.03
    a, _, _ = E.k_sample_topkp(logits, 1.0, top_p=0.75, seed=4, step=9)
    b, _, _ = E.k_sample_topkp(logits, 1.0, top_p=0.75, seed=4, step=9)
    c, _, _ = E.k_sample_topkp(logits, 1.0, top_p=0.75, seed=5, step=9)
    assert np.array_equal(a, b) and not np.array_equal(a, c)


def test_generation_with_filters():
    with E.Engine("tiny-llama", seed=1, max_batch=2, max_ctx=160) as eng:
        p = [list(range(10, 40)), list(range(10, 40)) + [7, 8, 9]]
        greedy = [g.tokens for g in eng.generate_ids(p, max_new_tokens=24)]
        # top_k = 1 and a vanishing nucleus are greedy whatever the temperature and seed (graph path included)
        assert [g.tokens for g in eng.generate_ids(p, max_new_tokens=24, temperature=0.9, seed=3, top_k=1)] == greedy
        assert [g.tokens for g in eng.generate_ids(p, max_new_tokens=24, temperature=1.5, seed=8, top_p=1e-6)] == greedy
        # a filter that keeps (all but one of) the tokens draws from the same noise stream as the unfiltered sampler
        plain = [g.tokens for g in eng.generate_ids(p, max_new_tokens=24, temperature=1.0, seed=5)]
        assert [g.tokens for g in eng.generate_ids(p, max_new_tokens=24, temperature=1.0, seed=5, top_k=eng.vocab - 1)] == plain
        a = [g.tokens for g in eng.generate_ids(p, max_new_tokens=24, temperature=1.0, seed=5, top_k=20, top_p=0.9)]
        b = [g.tokens for g in eng.generate_ids(p, max_new_tokens=24, temperature=1.0, seed=5, top_k=20, top_p=0.9)]
        c = [g.tokens for g in eng.generate_ids(p, max_new_tokens=24, temperature=1.0, seed=6, top_k=20, top_p=0.9)]
        assert a == b and a != c and a != plain and a != greedy
        # every drawn token is inside the top-20 of the logits the engine itself computes for that position
        eng.seq_reset()
        seq = p[0] + a[0]
        logits = eng.step_logits(seq, mode=0)
        for i, tok in enumerate(a[0]):
            row = logits[len(p[0]) - 1 + i]
            assert tok in np.argsort(-row)[:22]                     # 2 ranks of slack for batch-vs-single rounding
        with pytest.raises(E.EngineError, match="top_p"):
            eng.generate_ids(p, max_new_tokens=2, temperature=1.0, top_p=1.5)
