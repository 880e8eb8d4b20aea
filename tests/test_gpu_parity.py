"""End-to-end parity of the CUDA path against the CPU oracle, through the C-ABI.

Tolerances (north_star): per-step logits within 1e-2 relative (relative to max |logit| of the step);
greedy token sequence identical for the first 64 tokens -- with the margin-aware rule of SURVEY.md 8c:
a divergence is accepted only where the oracle's own top-1/top-2 margin is below the measured logit
error bound (a numerical tie), after which the comparison is teacher-forced."""
import os

import numpy as np
import pytest
import torch

from oracle.archs import ARCHS
from oracle.model import KVState, OracleModel
from theroundtaible_b200 import B200LocalLlmAdapter, AdapterError, engine as E

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TINY = ["tiny-llama", "tiny-llama-hd128", "tiny-qwen", "tiny-qwen-hd128"]
REL_TOL = 1e-2


def _rel_err(got, ref):
    return float(np.abs(got - ref).max() / np.abs(ref).max())


@pytest.mark.parametrize("name", TINY)
def test_teacher_forced_logits_match_oracle_and_hf(name):
    g = np.load(os.path.join(GOLD, f"hf_{name}.npz"))
    a, seed, ids = ARCHS[name], int(g["seed"]), g["ids"]
    ref = OracleModel(a, seed, max_pos=256).forward(ids, KVState(), round_kv=True, round_act=True).numpy()
    with E.Engine(name, seed=seed, max_batch=1, max_ctx=256) as eng:
        eng.seq_reset()
        got = np.concatenate([eng.step_logits(ids[:50]), eng.step_logits(ids[50:51]), eng.step_logits(ids[51:])])
    per_step = np.abs(got - ref).max(-1) / np.abs(ref).max(-1)
    assert per_step.max() <= REL_TOL, per_step.max()
    assert _rel_err(got, g["logits"]) <= REL_TOL                      # and within tolerance of HF fp32 itself
    assert np.median(per_step) <= 3e-3                                # typical error is far below the bar


@pytest.mark.parametrize("name", TINY)
def test_tensor_core_prefill_logits_match_oracle_and_decode_path(name):
    """mode 1 = tcgen05 GEMMs + flash prefill attention; mode 0 = decode kernels.  Same rounding points, so both must
    sit within the tolerance of the oracle and of each other; appending in several calls exercises chunk continuation."""
    g = np.load(os.path.join(GOLD, f"hf_{name}.npz"))
    a, seed, ids = ARCHS[name], int(g["seed"]), g["ids"]
    ref = OracleModel(a, seed, max_pos=256).forward(ids, KVState(), round_kv=True, round_act=True).numpy()
    with E.Engine(name, seed=seed, max_batch=1, max_ctx=256) as eng:
        eng.seq_reset()
        got1 = np.concatenate([eng.step_logits(ids[:37], mode=1), eng.step_logits(ids[37:38], mode=1), eng.step_logits(ids[38:], mode=1)])
        eng.seq_reset()
        got0 = eng.step_logits(ids, mode=0)
        eng.seq_reset()
        mixed = np.concatenate([eng.step_logits(ids[:50], mode=1), eng.step_logits(ids[50:], mode=0)])   # prefill then decode kernels
    scale = np.abs(ref).max(-1)
    assert (np.abs(got1 - ref).max(-1) / scale).max() <= REL_TOL
    assert (np.abs(got1 - got0).max(-1) / scale).max() <= REL_TOL
    assert (np.abs(mixed - ref).max(-1) / scale).max() <= REL_TOL
    assert np.median(np.abs(got1 - ref).max(-1) / scale) <= 3e-3


def test_prefill_paths_agree_on_generation():
    a = ARCHS["tiny-qwen-hd128"]
    rng = np.random.default_rng(77)
    common = rng.integers(0, a.vocab, 80).tolist()
    prompts = [common + rng.integers(0, a.vocab, 11 + 3 * i).tolist() for i in range(4)]
    oracle = OracleModel(a, 21, max_pos=256)
    with E.Engine("tiny-qwen-hd128", seed=21, max_batch=4, max_ctx=256) as eng:
        tc = eng.generate_ids(prompts, max_new_tokens=20)
    with E.Engine("tiny-qwen-hd128", seed=21, max_batch=4, max_ctx=256, flags=E.RT_FLAG_PREFILL_VIA_DECODE) as eng:
        slab = eng.generate_ids(prompts, max_new_tokens=20)
    for p, x, y in zip(prompts, tc, slab):
        _greedy_compare(x.tokens, oracle, p, 20, err_bound=REL_TOL * 2.5)
        _greedy_compare(y.tokens, oracle, p, 20, err_bound=REL_TOL * 2.5)
    assert sum(x.tokens == y.tokens for x, y in zip(tc, slab)) >= 3


def _greedy_compare(eng_tokens, oracle: OracleModel, prompt, n_new, err_bound):
    """Margin-aware greedy comparison.  Returns number of accepted tie-breaks."""
    kv = KVState()
    logits = oracle.forward(prompt, kv, round_act=True, last_only=True)[-1].numpy()
    ties = 0
    for i in range(n_new):
        top2 = np.partition(logits, -2)[-2:]
        margin = float(top2[1] - top2[0])
        want = int(logits.argmax())
        if eng_tokens[i] != want:
            assert margin <= err_bound, f"step {i}: engine {eng_tokens[i]} != oracle {want}, margin {margin} > bound {err_bound}"
            ties += 1
        logits = oracle.forward([eng_tokens[i]], kv, round_act=True, last_only=True)[-1].numpy()   # teacher-force
    return ties


@pytest.mark.parametrize("name,plen", [("tiny-llama", 40), ("tiny-qwen", 33), ("tiny-llama-hd128", 70), ("tiny-qwen-hd128", 17)])
def test_greedy_64_tokens_match_oracle(name, plen):
    a = ARCHS[name]
    oracle = OracleModel(a, 3, max_pos=256)
    prompt = np.random.default_rng(plen).integers(0, a.vocab, plen).tolist()
    with E.Engine(name, seed=3, max_batch=2, max_ctx=256) as eng:
        gen = eng.generate_ids([prompt], max_new_tokens=64)[0]
        nog = E.Engine(name, seed=3, max_batch=2, max_ctx=256, flags=E.RT_FLAG_NO_GRAPH)
        gen2 = nog.generate_ids([prompt], max_new_tokens=64)[0]
        nog.close()
    assert len(gen.tokens) == 64 and gen.n_prompt_tokens == plen and gen.decode_steps == 63
    assert gen.tokens == gen2.tokens                                  # CUDA-graph replay == kernel-by-kernel
    ties = _greedy_compare(gen.tokens, oracle, prompt, 64, err_bound=REL_TOL * 2.5)
    assert ties <= 2


@pytest.mark.parametrize("name,n,shared,own", [("tiny-llama", 3, 64, 9), ("tiny-qwen-hd128", 5, 48, 20),
                                                ("tiny-llama-hd128", 8, 100, 5), ("tiny-qwen", 2, 0, 30)])
def test_batch_with_shared_prefix_equals_unshared_and_oracle(name, n, shared, own):
    a = ARCHS[name]
    rng = np.random.default_rng(n * 100 + shared)
    common = rng.integers(0, a.vocab, shared).tolist()
    prompts = [common + rng.integers(0, a.vocab, own + i).tolist() for i in range(n)]     # ragged suffixes
    with E.Engine(name, seed=9, max_batch=8, max_ctx=256) as eng:
        gens = eng.generate_ids(prompts, max_new_tokens=24)
    with E.Engine(name, seed=9, max_batch=8, max_ctx=256, flags=E.RT_FLAG_NO_PREFIX_SHARE) as eng:
        plain = eng.generate_ids(prompts, max_new_tokens=24)
    oracle = OracleModel(a, 9, max_pos=256)
    for i in range(n):
        assert gens[i].prefix_shared_tokens == (0 if i == 0 else (shared // 16) * 16)
        assert plain[i].prefix_shared_tokens == 0
        _greedy_compare(gens[i].tokens, oracle, prompts[i], 24, err_bound=REL_TOL * 2.5)
        _greedy_compare(plain[i].tokens, oracle, prompts[i], 24, err_bound=REL_TOL * 2.5)
    # cascade vs plain differ by fp32 merge order only, so they part ways only at near-ties -- which the margin-aware
    # comparisons above already police for both; on top of that most knights must come out token-identical
    same = sum(g.tokens == p.tokens for g, p in zip(gens, plain))
    assert same >= (n + 1) // 2


def test_real_architecture_truncated_depth_llama32_1b():
    # true Llama-3.2-1B dimensions, 2 layers (SURVEY.md 8c "large configs": truncated depth, stated)
    a = ARCHS["llama-3.2-1b"].with_layers(2)
    oracle = OracleModel(a, 0, max_pos=512)
    rng = np.random.default_rng(1234 + 2)
    ids = rng.integers(0, a.vocab, 160)
    ref = oracle.forward(ids, KVState(), round_act=True).numpy()
    with E.Engine("llama-3.2-1b@L2", seed=0, max_batch=3, max_ctx=512) as eng:
        eng.seq_reset()
        got = eng.step_logits(ids)
        per_step = np.abs(got - ref).max(-1) / np.abs(ref).max(-1)
        assert per_step.max() <= REL_TOL
        prompts = [ids[:128].tolist() + rng.integers(0, a.vocab, 16).tolist() for _ in range(3)]
        gens = eng.generate_ids(prompts, max_new_tokens=32)
        for p, g in zip(prompts, gens):
            _greedy_compare(g.tokens, oracle, p, 32, err_bound=REL_TOL * float(np.abs(ref).max()))
            assert g.prefix_shared_tokens in (0, 128)


def test_config1_architecture_full_depth_qwen25_05b():
    """BASELINE.json configs[0] names Qwen2.5-0.5B: the FULL 24-layer architecture (qkv bias, 14/2 heads, tied head,
    151,936-token vocab) against the oracle: teacher-forced logits through both kernel paths and 64 greedy tokens for
    3 knights sharing a prefix (margin-aware rule)."""
    a = ARCHS["qwen2.5-0.5b"]
    oracle = OracleModel(a, 0, max_pos=512)
    rng = np.random.default_rng(1234 + 1)
    ids = rng.integers(0, a.vocab, 96)
    ref = oracle.forward(ids, KVState(), round_act=True).numpy()
    scale = np.abs(ref).max(-1)
    with E.Engine("qwen2.5-0.5b", seed=0, max_batch=3, max_ctx=512) as eng:
        eng.seq_reset()
        got1 = eng.step_logits(ids, mode=1)
        eng.seq_reset()
        got0 = eng.step_logits(ids, mode=0)
        assert (np.abs(got1 - ref).max(-1) / scale).max() <= REL_TOL
        assert (np.abs(got0 - ref).max(-1) / scale).max() <= REL_TOL
        common = ids[:64].tolist()
        prompts = [common + rng.integers(0, a.vocab, 20 + i).tolist() for i in range(3)]
        gens = eng.generate_ids(prompts, max_new_tokens=64)
    bound = REL_TOL * float(np.abs(ref).max())
    ties = sum(_greedy_compare(g.tokens, oracle, p, 64, err_bound=bound) for p, g in zip(prompts, gens))
    assert ties <= 6


def test_real_architecture_truncated_depth_llama31_8b():
    """True Llama-3.1-8B dimensions (head_dim 128, 32/8 heads, hidden 4096, FFN 14336, untied 128k head) at 2 layers
    (truncated depth, stated -- SURVEY.md 8c): both kernel paths against the oracle, then 5 knights with a shared prefix."""
    a = ARCHS["llama-3.1-8b"].with_layers(2)
    oracle = OracleModel(a, 0, max_pos=512)
    rng = np.random.default_rng(1234 + 3)
    ids = rng.integers(0, a.vocab, 200)
    ref = oracle.forward(ids, KVState(), round_act=True).numpy()
    scale = np.abs(ref).max(-1)
    with E.Engine("llama-3.1-8b@L2", seed=0, max_batch=5, max_ctx=512) as eng:
        eng.seq_reset()
        got1 = eng.step_logits(ids, mode=1)
        eng.seq_reset()
        got0 = eng.step_logits(ids[:70], mode=0)
        assert (np.abs(got1 - ref).max(-1) / scale).max() <= REL_TOL
        assert (np.abs(got0 - ref[:70]).max(-1) / scale[:70]).max() <= REL_TOL
        prompts = [ids[:160].tolist() + rng.integers(0, a.vocab, 12 + i).tolist() for i in range(5)]
        gens = eng.generate_ids(prompts, max_new_tokens=24)
    bound = REL_TOL * float(np.abs(ref).max())
    for p, g in zip(prompts, gens):
        _greedy_compare(g.tokens, oracle, p, 24, err_bound=bound)
        assert g.prefix_shared_tokens in (0, 160)


def test_cross_round_prefix_cache_reuses_a_knights_own_prompt():
    """in the real flow knights speak one at a time (SEQ) and a knight's next-round prompt extends
    its previous one (src/utils/prompt.ts:60-77).  With RT_FLAG_PREFIX_CACHE the engine keeps the prompt's KV pages and
    prefills only the new tail; results must still match the oracle, and a small pool must evict instead of failing."""
    name = "tiny-llama-hd128"
    a = ARCHS[name]
    oracle = OracleModel(a, 6, max_pos=512)
    rng = np.random.default_rng(8)
    heads = [rng.integers(0, a.vocab, 70 + 5 * k).tolist() for k in range(3)]      # three knights, distinct prompt heads
    with E.Engine(name, seed=6, max_batch=1, max_ctx=512, flags=E.RT_FLAG_PREFIX_CACHE) as eng:
        round1 = [eng.generate_ids([h], max_new_tokens=12)[0] for h in heads]
        assert all(g.prefix_cached_tokens == 0 for g in round1)
        prompts2 = [h + rng.integers(0, a.vocab, 40).tolist() for h in heads]      # transcript grew
        round2 = [eng.generate_ids([p], max_new_tokens=12)[0] for p in prompts2]
        for h, p, g in zip(heads, prompts2, round2):
            assert g.prefix_cached_tokens == (len(h) // 16) * 16                   # whole pages of the old prompt
            _greedy_compare(g.tokens, oracle, p, 12, err_bound=REL_TOL * 2.5)
        again = eng.generate_ids([prompts2[0]], max_new_tokens=12)[0]              # identical prompt: all but the last page part
        assert again.prefix_cached_tokens == ((len(prompts2[0]) - 1) // 16) * 16
        _greedy_compare(again.tokens, oracle, prompts2[0], 12, err_bound=REL_TOL * 2.5)
        # batch call: knight 0's pages come from the cache, the others share / prefill as usual
        common = prompts2[0][:96]
        batch = [common + rng.integers(0, a.vocab, 10 + i).tolist() for i in range(1)]
        g = eng.generate_ids(batch, max_new_tokens=8)[0]
        assert g.prefix_cached_tokens == 96
        _greedy_compare(g.tokens, oracle, batch[0], 8, err_bound=REL_TOL * 2.5)
    # pool of 40 pages (640 tokens): the third distinct 300-token prompt can only fit if older entries are evicted
    kv_bytes = 40 * 2 * a.layers * a.n_kv * 16 * a.head_dim * 2
    with E.Engine(name, seed=6, max_batch=1, max_ctx=512, flags=E.RT_FLAG_PREFIX_CACHE, kv_pool_bytes=kv_bytes) as eng:
        for k in range(4):
            p = rng.integers(0, a.vocab, 300).tolist()
            g = eng.generate_ids([p], max_new_tokens=8)[0]
            assert len(g.tokens) == 8 and g.prefix_cached_tokens == 0


def test_temperature_sampling_is_seeded_and_varies():
    with E.Engine("tiny-llama", seed=1, max_batch=2, max_ctx=128) as eng:
        p = list(range(10, 40))
        a = eng.generate_ids([p], max_new_tokens=20, temperature=1.0, seed=5)[0].tokens
        b = eng.generate_ids([p], max_new_tokens=20, temperature=1.0, seed=5)[0].tokens
        c = eng.generate_ids([p], max_new_tokens=20, temperature=1.0, seed=6)[0].tokens
        g = eng.generate_ids([p], max_new_tokens=20)[0].tokens
    assert a == b and a != c and a != g


def test_errors_cross_the_boundary_as_classifiable_text():
    with E.Engine("tiny-llama", seed=1, max_batch=2, max_ctx=64) as eng:
        with pytest.raises(E.EngineError, match="context length exceeded"):
            eng.generate_ids([list(range(60))], max_new_tokens=32)
        with pytest.raises(E.EngineError, match="token id out of range"):
            eng.generate_ids([[999999]], max_new_tokens=2)
        with pytest.raises(E.EngineError, match="batch size"):
            eng.generate_ids([[1]] * 3, max_new_tokens=2)
        # still usable after errors
        assert len(eng.generate_ids([[1, 2, 3]], max_new_tokens=4)[0].tokens) == 4
    eng = E.Engine("llama-3.2-1b", seed=0, max_batch=1, max_ctx=4096)
    ad = B200LocalLlmAdapter("b200://0", "llama-3.2-1b", "Lancelot", max_new_tokens=3000, ignore_eos=True, engine=eng)
    with pytest.raises(AdapterError) as ei:
        ad.execute("x" * 200, timeout_ms=20)                          # 20 ms budget: must time out
    assert ei.value.kind == "timeout" and "timed out" in str(ei.value)
    eng.close()


def test_adapter_execute_returns_nonempty_text_and_round_batches():
    ad = B200LocalLlmAdapter("b200://0", "tiny-qwen", "Merlin", max_ctx=2048, max_new_tokens=40)
    assert ad.is_available()
    text = ad.execute("Onderwerp: moeten we de cache herschrijven?")
    assert isinstance(text, str) and len(text) > 0
    assert ad.get_max_source_chars() == 2000
    head = "Je neemt deel aan een TheRoundtAIble discussie.\nREGELS: " + "x" * 200 + "\nJe naam is: "
    prompts = [head + n for n in ("Claude", "Gemini", "GPT")]
    texts = ad.execute_round(prompts)
    assert len(texts) == 3 and all(len(t) > 0 for t in texts)
    batch = [g.tokens for g in ad.last_round]
    assert ad.execute(prompts[0]) == E.detokenize(ad.last_generation.tokens)
    seq = ad.last_generation.tokens
    # SEQ (one knight per call) and BATCH (one call per round) both follow the oracle under the margin rule: they can
    # only part ways at a numerical tie of the oracle's own top two logits
    oracle = OracleModel(ARCHS["tiny-qwen"], 0, max_pos=2048)
    for p, toks in [(prompts[0], seq)] + list(zip(prompts, batch)):
        _greedy_compare(toks, oracle, E.tokenize_chat(p), len(toks), err_bound=REL_TOL * 2.5)


def test_config2_architecture_full_depth_llama32_1b():
    """BASELINE.json configs[1] names Llama-3.2-1B: the FULL 16-layer architecture (llama3 rope scaling, 32/8 heads, tied
    128,256-token head) against the oracle: teacher-forced logits through both kernel paths, then 3 knights sharing a
    prefix, greedy, under the margin-aware rule.

    The bar is north_star's: every step's logits within 1e-2 of max |logit|, in BOTH kernel paths, and 64 greedy tokens
    under the margin rule.  Round 1 missed it at full depth (1.0-1.3e-2) because bf16 projection inputs put the
    rounding-noise floor of the contract itself at 1.2e-2 for this model (scripts/noise_floor_probe.py); projection
    inputs are fp16 since round 2 (floor 2.5e-3), so the bar holds with margin."""
    a = ARCHS["llama-3.2-1b"]
    oracle = OracleModel(a, 0, max_pos=256)
    rng = np.random.default_rng(1234 + 2)
    ids = rng.integers(0, a.vocab, 72)
    ref = oracle.forward(ids, KVState(), round_act=True).numpy()
    scale = np.abs(ref).max(-1)
    with E.Engine("llama-3.2-1b", seed=0, max_batch=3, max_ctx=256) as eng:
        eng.seq_reset()
        got1 = eng.step_logits(ids, mode=1)
        eng.seq_reset()
        got0 = eng.step_logits(ids[:40], mode=0)
        for got, want, sc in ((got1, ref, scale), (got0, ref[:40], scale[:40])):
            worst = (np.abs(got - want).max(-1) / sc).max()
            rms = (np.sqrt(((got - want) ** 2).mean(-1)) / sc).max()
            assert worst <= REL_TOL, worst
            assert rms <= 1e-3, rms
        prompts = [ids[:48].tolist() + rng.integers(0, a.vocab, 10 + i).tolist() for i in range(3)]
        gens = eng.generate_ids(prompts, max_new_tokens=64)
    bound = REL_TOL * float(np.abs(ref).max())
    ties = sum(_greedy_compare(g.tokens, oracle, p, 64, err_bound=bound) for p, g in zip(prompts, gens))
    assert ties <= 4


@pytest.mark.parametrize("model,layers,n_ids", [("llama-3.1-8b", 8, 96), ("qwen2.5-32b", 1, 64), ("llama-3.1-70b", 1, 64)])
def test_large_configs_true_dimensions_truncated_depth(model, layers, n_ids):
    """SURVEY.md 8c "Large configs": the benchmarked models the CPU oracle cannot hold at full depth are checked at their
    TRUE dimensions with the layer count truncated (stated here and in every report): Llama-3.1-8B at 8 of 32 layers,
    Qwen2.5-32B (hidden 5,120, 40/8 heads, qkv bias, FFN 27,648, 152k vocab) and Llama-3.1-70B (hidden 8,192, 64/8 heads,
    FFN 28,672) at one decoder layer each.  Both kernel paths against the oracle at the north_star bar."""
    a = ARCHS[model].with_layers(layers)
    oracle = OracleModel(a, 0, max_pos=256)
    ids = np.random.default_rng(layers + n_ids).integers(0, a.vocab, n_ids)
    ref = oracle.forward(ids, KVState(), round_act=True).numpy()
    del oracle
    scale = np.abs(ref).max(-1)
    with E.Engine(f"{model}@L{layers}", seed=0, max_batch=1, max_ctx=256) as eng:
        eng.seq_reset()
        got1 = eng.step_logits(ids, mode=1)
        eng.seq_reset()
        got0 = eng.step_logits(ids[:40], mode=0)
    assert (np.abs(got1 - ref).max(-1) / scale).max() <= REL_TOL
    assert (np.abs(got0 - ref[:40]).max(-1) / scale[:40]).max() <= REL_TOL


def test_config2_full_size_properties():
    """BASELINE config 2 at its full size (3 knights, Llama-3.2-1B, 1,024 shared + 64 own prompt tokens) is beyond what
    the CPU oracle finishes in seconds, so it is checked through properties that do not need it: a round is
    bit-reproducible; a sequence's tokens do not depend on which other knights share its batch or on whether its prefix
    came from the shared pages, the cross-call cache or its own prefill (up to one near-tie); every generated id is a
    valid token and the counts add up."""
    w = dict(knights=3, shared=1024, own=64, new=48)
    rng = np.random.default_rng(1234 + 2)
    with E.Engine("llama-3.2-1b", seed=0, max_batch=3, max_ctx=2048) as eng:
        vocab = eng.vocab
        common = rng.integers(0, vocab, w["shared"]).tolist()
        prompts = [common + rng.integers(0, vocab, w["own"]).tolist() for _ in range(w["knights"])]
        a = eng.generate_ids(prompts, max_new_tokens=w["new"])
        b = eng.generate_ids(prompts, max_new_tokens=w["new"])
        assert [g.tokens for g in a] == [g.tokens for g in b]                       # run-to-run bit-identical
        assert all(len(g.tokens) == w["new"] and all(0 <= t < vocab for t in g.tokens) for g in a)
        assert [g.prefix_shared_tokens for g in a] == [0, 1024, 1024] and a[0].n_prompt_tokens == 1088
        solo = [eng.generate_ids([p], max_new_tokens=w["new"])[0] for p in prompts]        # SEQ: one knight per call
    same = sum(x.tokens == y.tokens for x, y in zip(a, solo))
    assert same >= 2, [next((i for i, (s, t) in enumerate(zip(x.tokens, y.tokens)) if s != t), None) for x, y in zip(a, solo)]
    with E.Engine("llama-3.2-1b", seed=0, max_batch=3, max_ctx=2048, flags=E.RT_FLAG_PREFIX_CACHE) as eng:
        first = eng.generate_ids([prompts[0]], max_new_tokens=w["new"])[0]
        again = eng.generate_ids([prompts[0]], max_new_tokens=w["new"])[0]
        other = eng.generate_ids([prompts[1]], max_new_tokens=w["new"])[0]
    assert first.prefix_cached_tokens == 0 and again.prefix_cached_tokens >= 1072 and other.prefix_cached_tokens == 1024
    assert first.tokens == solo[0].tokens                                           # the cache flag alone changes nothing
    # KV from the cache == KV recomputed; the 16-token tail of the second call runs through a different kernel path than
    # the full prompt did, so a late near-tie may part them
    assert again.tokens == first.tokens or again.tokens[:8] == first.tokens[:8]
    assert sum(x == y for x, y in zip(other.tokens, solo[1].tokens)) >= w["new"] - 1 or other.tokens[:8] == solo[1].tokens[:8]


def test_opt_in_decode_variants_reproduce_the_default_tokens(monkeypatch):
    """Switchable variants of the decode step that are kept for measurement (profiles/r1_summary.md) must not change
    results: several steps per CUDA-graph launch (with the late QKV release) replay the same kernels in the same order."""
    a = ARCHS["tiny-llama-hd128"]
    rng = np.random.default_rng(5)
    common = rng.integers(0, a.vocab, 70).tolist()
    prompts = [common + rng.integers(0, a.vocab, 7 + i).tolist() for i in range(3)]

    def run():
        with E.Engine("tiny-llama-hd128", seed=2, max_batch=3, max_ctx=256) as eng:
            return [g.tokens for g in eng.generate_ids(prompts, max_new_tokens=40)]
    base = run()
    monkeypatch.setenv("RT_GRAPH_STEPS", "8")
    assert run() == base                                       # same kernels, same order: bit-identical
    monkeypatch.delenv("RT_GRAPH_STEPS")
    oracle = OracleModel(a, 2, max_pos=256)
    for p, t in zip(prompts, base):
        _greedy_compare(t, oracle, p, 40, err_bound=REL_TOL * 2.5)


def test_tcgen05_prefill_attention_head_dim_64(monkeypatch):
    """The head_dim-64 instantiation of the tcgen05 prefill attention (RT_PFA_TC5=2; not the default: no measurable gain)
    on true Llama-3.2-1B dimensions at 2 layers, against the oracle."""
    a = ARCHS["llama-3.2-1b"].with_layers(2)
    ids = np.random.default_rng(7).integers(0, a.vocab, 200)
    ref = OracleModel(a, 0, max_pos=512).forward(ids, KVState(), round_act=True).numpy()
    monkeypatch.setenv("RT_PFA_TC5", "2")
    with E.Engine("llama-3.2-1b@L2", seed=0, max_batch=1, max_ctx=512) as eng:
        eng.seq_reset()
        got = np.concatenate([eng.step_logits(ids[:150], mode=1), eng.step_logits(ids[150:], mode=1)])
    assert (np.abs(got - ref).max(-1) / np.abs(ref).max(-1)).max() <= REL_TOL


@pytest.mark.parametrize("name", ["tiny-llama-hd128", "tiny-qwen-hd128"])
def test_tcgen05_prefill_attention_matches_oracle_and_the_mma_sync_kernel(name, monkeypatch):
    """Head_dim 128 takes the tcgen05 prefill attention by default (RT_PFA_TC5=0 selects the mma.sync kernel) (attn_prefill_tc5.cuh: S and O in TMEM, K/V pages by TMA from the
    paged pool, G = 4 and G = 5 head groups -> 128 and 125 query rows per CTA).  Same bar as the mma.sync kernel, in one
    call and in a continued sequence (second call attends to pages the first one wrote)."""
    g = np.load(os.path.join(GOLD, f"hf_{name}.npz"))
    a, seed, ids = ARCHS[name], int(g["seed"]), g["ids"]
    ref = OracleModel(a, seed, max_pos=256).forward(ids, KVState(), round_kv=True, round_act=True).numpy()
    scale = np.abs(ref).max(-1)
    monkeypatch.setenv("RT_PFA_TC5", "0")
    with E.Engine(name, seed=seed, max_batch=1, max_ctx=256) as eng:
        eng.seq_reset()
        base = eng.step_logits(ids, mode=1)
    monkeypatch.setenv("RT_PFA_TC5", "2")
    with E.Engine(name, seed=seed, max_batch=1, max_ctx=256) as eng:
        eng.seq_reset()
        one = eng.step_logits(ids, mode=1)
        eng.seq_reset()
        parts = np.concatenate([eng.step_logits(ids[:33], mode=1), eng.step_logits(ids[33:], mode=1)])
    for got in (one, parts):
        assert (np.abs(got - ref).max(-1) / scale).max() <= REL_TOL
        assert (np.abs(got - base).max(-1) / scale).max() <= REL_TOL
        assert np.median(np.abs(got - ref).max(-1) / scale) <= 3e-3


@pytest.mark.parametrize("name", ["tiny-llama-hd128", "llama-3.1-8b@L2"])
def test_decode_projections_over_fp16_twins_5_to_8_slots(name, monkeypatch):
    """Decode steps with 5..8 token slots read the fp16 twins of the weights (gemv.cuh mode 2: one f16 x f16 MMA per
    fragment instead of bf16 x (hi + lo)): the same products, so the logits must sit within the tolerance of the oracle AND
    within rounding of the bf16-weight path (RT_DEC_F16=0)."""
    base, _, lay = name.partition("@L")
    a = ARCHS[base].with_layers(int(lay)) if lay else ARCHS[base]
    oracle = OracleModel(a, 4, max_pos=128)
    rng = np.random.default_rng(55)
    ids = rng.integers(0, a.vocab, 5 + 6 + 7 + 8)
    ref = oracle.forward(ids, KVState(), round_act=True).numpy()
    scale = np.abs(ref).max(-1)

    def run():
        with E.Engine(name, seed=4, max_batch=8, max_ctx=128) as eng:
            eng.seq_reset()
            out, at = [], 0
            for n in (5, 6, 7, 8):                       # one decode step of n slots per call
                out.append(eng.step_logits(ids[at:at + n], mode=0))
                at += n
            prompts = [ids[:12].tolist() + [int(t)] for t in ids[12:18]]       # 6 knights -> 6 slots per decode step
            toks = [g.tokens for g in eng.generate_ids(prompts, max_new_tokens=16)]
        return np.concatenate(out), prompts, toks

    twin, prompts, toks = run()
    monkeypatch.setenv("RT_DEC_F16", "0")
    plain, _, toks_plain = run()
    assert (np.abs(twin - ref).max(-1) / scale).max() <= REL_TOL
    assert (np.abs(plain - ref).max(-1) / scale).max() <= REL_TOL
    # not bit-equal: the accumulation order differs, and a last-bit difference that flips the fp16 rounding of an activation or
    # the bf16 rounding of a K/V element travels on (measured 1.3e-3 at true 8B dimensions; the contract's own noise floor
    # between two evaluation orders is 2.5e-3, scripts/noise_floor_probe.py)
    assert (np.abs(twin - plain).max(-1) / scale).max() <= 3e-3
    bound = REL_TOL * float(np.abs(ref).max())
    for p, t in zip(prompts, toks):
        _greedy_compare(t, oracle, p, 16, err_bound=bound)
    assert sum(x == y for x, y in zip(toks, toks_plain)) >= 4
