"""Edge cases of the generation path through the C-ABI: empty and one-token prompts, ragged batches of the maximum size,
a prompt that fills the context exactly, one new token, prefixes shorter than a KV page, contexts on both sides of the
switch between the two decode-attention kernels."""
import numpy as np
import pytest

from oracle.archs import ARCHS
from oracle.model import OracleModel
from tests.test_gpu_parity import REL_TOL, _greedy_compare
from theroundtaible_b200 import B200LocalLlmAdapter, engine as E

pytestmark = pytest.mark.gpu


def test_empty_and_one_token_prompts():
    with E.Engine("tiny-llama", seed=1, max_batch=2, max_ctx=128) as eng:
        g = eng.generate("", max_new_tokens=5, ignore_eos=True)             # chat wrapper only: <bos><user><assistant>
        assert g.n_prompt_tokens == 3 and len(g.tokens) == 5 and g.text
        one = eng.generate_ids([[7]], max_new_tokens=6)[0]
        assert one.n_prompt_tokens == 1 and len(one.tokens) == 6
        oracle = OracleModel(ARCHS["tiny-llama"], 1, max_pos=64)
        _greedy_compare(one.tokens, oracle, [7], 6, err_bound=REL_TOL * 2.5)
        with pytest.raises(E.EngineError):
            eng.generate_ids([[]], max_new_tokens=2)                          # an empty id list has nothing to attend to


def test_ragged_batch_of_eight_and_single_new_token():
    a = ARCHS["tiny-qwen"]
    rng = np.random.default_rng(8)
    prompts = [rng.integers(0, a.vocab, n).tolist() for n in (1, 2, 15, 16, 17, 31, 64, 100)]   # around page boundaries
    oracle = OracleModel(a, 6, max_pos=256)
    with E.Engine("tiny-qwen", seed=6, max_batch=8, max_ctx=160) as eng:
        gens = eng.generate_ids(prompts, max_new_tokens=12)
        first = eng.generate_ids(prompts, max_new_tokens=1)
    for p, g, f in zip(prompts, gens, first):
        assert g.n_prompt_tokens == len(p) and len(g.tokens) == 12 and g.prefix_shared_tokens == 0
        assert f.tokens == g.tokens[:1] and f.decode_steps == 0
        _greedy_compare(g.tokens, oracle, p, 12, err_bound=REL_TOL * 2.5)


def test_prompt_that_fills_the_context_exactly():
    a = ARCHS["tiny-llama"]
    prompt = np.random.default_rng(3).integers(0, a.vocab, 120).tolist()
    with E.Engine("tiny-llama", seed=1, max_batch=1, max_ctx=128) as eng:
        g = eng.generate_ids([prompt], max_new_tokens=8)[0]                   # 120 + 8 = max_ctx
        assert len(g.tokens) == 8
        with pytest.raises(E.EngineError, match="context length exceeded"):
            eng.generate_ids([prompt], max_new_tokens=9)
    oracle = OracleModel(a, 1, max_pos=128)
    _greedy_compare(g.tokens, oracle, prompt, 8, err_bound=REL_TOL * 2.5)


def test_shared_prefix_shorter_than_a_page_and_identical_prompts():
    a = ARCHS["tiny-llama"]
    rng = np.random.default_rng(4)
    common = rng.integers(0, a.vocab, 9).tolist()                             # < 16 tokens: nothing page-aligned to share
    prompts = [common + rng.integers(0, a.vocab, 5).tolist() for _ in range(3)]
    same = [common + [1, 2, 3]] * 3                                           # identical prompts: everything but the tail shared
    with E.Engine("tiny-llama", seed=1, max_batch=3, max_ctx=128) as eng:
        g1 = eng.generate_ids(prompts, max_new_tokens=10)
        g2 = eng.generate_ids(same, max_new_tokens=10)
    assert [g.prefix_shared_tokens for g in g1] == [0, 0, 0]
    assert g2[0].tokens == g2[1].tokens == g2[2].tokens
    oracle = OracleModel(a, 1, max_pos=128)
    for p, g in zip(prompts, g1):
        _greedy_compare(g.tokens, oracle, p, 10, err_bound=REL_TOL * 2.5)


@pytest.mark.parametrize("plen", [4000, 4200])
def test_both_sides_of_the_attention_kernel_switch(plen):
    """Contexts up to 4,096 tokens take the cluster-merged attention kernel, longer ones the split-KV kernel + merge; a
    generation that crosses the limit keeps the kernel chosen for its longest context.  Both must match the oracle."""
    a = ARCHS["tiny-llama-hd128"]
    prompt = np.random.default_rng(plen).integers(0, a.vocab, plen).tolist()
    oracle = OracleModel(a, 2, max_pos=plen + 32)
    with E.Engine("tiny-llama-hd128", seed=2, max_batch=1, max_ctx=plen + 32) as eng:
        g = eng.generate_ids([prompt], max_new_tokens=12)[0]
    _greedy_compare(g.tokens, oracle, prompt, 12, err_bound=REL_TOL * 2.5)


def test_adapter_timeout_and_reuse_after_error():
    ad = B200LocalLlmAdapter("b200://0", "tiny-llama", "Kay", max_ctx=256, max_new_tokens=200, ignore_eos=True)
    with pytest.raises(Exception) as ei:
        ad.execute("x" * 400)                                                 # 403 prompt tokens > 256
    assert "context length exceeded" in str(ei.value)
    assert ad.execute("short prompt")                                          # the engine is still usable
    assert ad.last_generation.n_prompt_tokens == len("short prompt") + 3


def test_eos_ids_come_from_the_configuration():
    """rt_config.eos_ids (filled from a checkpoint's config.json by config_from_hf) decide where generation stops; the
    literal id 2 is only the default of the byte tokenizer."""
    prompt = list(range(20, 60))
    with E.Engine("tiny-llama", seed=3, max_batch=1, max_ctx=128) as eng:
        free = eng.generate_ids([prompt], max_new_tokens=12)[0].tokens
    stop_at = free.index(free[-1])                     # first occurrence of the id that will serve as <eos>
    cfg = E.arch_config("tiny-llama")
    cfg.n_eos, cfg.eos_ids[0], cfg.eos_ids[1] = 2, (free[-1] + 1) % 512 if (free[-1] + 1) % 512 not in free else free[-1], free[-1]
    with E.Engine(cfg, seed=3, max_batch=1, max_ctx=128) as eng:
        g = eng.generate_ids([prompt], max_new_tokens=12, ignore_eos=False)[0]
        assert g.tokens == free[:stop_at] and g.finish_reason == 1
        assert eng.generate_ids([prompt], max_new_tokens=12, ignore_eos=True)[0].tokens == free


@pytest.mark.parametrize("name,knights,shared,own", [("tiny-llama-hd128", 3, 4300, 30), ("tiny-qwen-hd128", 2, 4200, 17),
                                                      ("tiny-qwen-hd128", 1, 0, 4500), ("llama-3.2-1b@L2", 5, 4150, 40)])
def test_tcgen05_decode_attention_long_contexts(name, knights, shared, own, monkeypatch):
    """Contexts beyond the cluster kernel's 4,096 tokens take the tcgen05 decode attention (attn_decode_tc5.cuh): query rows
    = (slot, q-head) pairs in TMEM, one CTA per key range, shared-prefix ranges serving every knight at once (G = 4 and
    G = 5 groups, head_dim 128 and 64, with and without a shared prefix).  Must match the oracle under the margin rule and
    the mma.sync split-KV kernel it replaces (RT_DEC5=0) up to near-ties."""
    base = name.split("@")[0]
    a = ARCHS[base].with_layers(2) if "@L2" in name else ARCHS[base]
    rng = np.random.default_rng(shared + own)
    common = rng.integers(0, a.vocab, shared).tolist()
    prompts = [common + rng.integers(0, a.vocab, own + 3 * i).tolist() for i in range(knights)]
    ctx = shared + own + 64
    monkeypatch.setenv("RT_DEC5", "2")             # force it also where the engine's heuristic keeps the mma.sync kernel
    with E.Engine(name, seed=5, max_batch=knights, max_ctx=ctx) as eng:
        new = eng.generate_ids(prompts, max_new_tokens=10)
    monkeypatch.setenv("RT_DEC5", "0")
    with E.Engine(name, seed=5, max_batch=knights, max_ctx=ctx) as eng:
        old = eng.generate_ids(prompts, max_new_tokens=10)
    oracle = OracleModel(a, 5, max_pos=ctx)
    bound = REL_TOL * 2.5
    for p, g, o in zip(prompts, new, old):
        _greedy_compare(g.tokens, oracle, p, 10, err_bound=bound)
        _greedy_compare(o.tokens, oracle, p, 10, err_bound=bound)
    assert sum(g.tokens == o.tokens for g, o in zip(new, old)) >= (knights + 1) // 2
