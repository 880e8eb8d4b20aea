"""The restated turn loop (theroundtaible_b200/roundtable.py) driving the real adapter on the GPU: SEQ mode on the
reference's prompt layout, with the cross-call prefix cache doing what SURVEY.md 8f row 2 asks of it."""
import random

import pytest

pytestmark = pytest.mark.gpu


def test_two_round_discussion_reuses_each_knights_previous_prompt():
    from theroundtaible_b200 import B200LocalLlmAdapter, Engine
    from theroundtaible_b200 import roundtable as R
    from theroundtaible_b200.engine import RT_FLAG_PREFIX_CACHE, tokenize_chat
    knights = [R.KnightConfig(n, f"local-llm-{i}", ["analysis"], i) for i, n in enumerate(["Merlin", "Gawain", "Percival"])]
    with Engine("tiny-llama", seed=3, max_batch=2, max_ctx=8192, kv_pool_bytes=64 << 20, flags=RT_FLAG_PREFIX_CACHE) as eng:
        adapters = {k.adapter: B200LocalLlmAdapter("b200://0", "tiny-llama", k.name, max_new_tokens=24, ignore_eos=True, engine=eng)
                    for k in knights}
        seen = []

        def on_turn(turn, adapter):
            g = adapter.last_generation
            seen.append((turn.knight, turn.round, g.n_prompt_tokens, g.prefix_cached_tokens))
        res = R.run_discussion(adapters, knights, "Where should the cache live?", max_rounds=2, rng=random.Random(3),
                               on_turn=on_turn)
    assert len(res.all_rounds) == 6 and not any(t.error for t in res.turns) and not res.consensus
    assert all(e.response for e in res.all_rounds)                       # the adapter rejects empty output
    # every later prompt carries every earlier response verbatim
    prompts = {}
    all_rounds = []
    for e in res.all_rounds:
        k = next(k for k in knights if k.name == e.knight)
        p = R.build_full_prompt(R.build_system_prompt(k, knights, "Where should the cache live?", "", all_rounds),
                                "Where should the cache live?")
        assert len(tokenize_chat(p)) == next(s[2] for s in seen if s[0] == e.knight and s[1] == e.round)
        prompts[(e.knight, e.round)] = tokenize_chat(p)
        all_rounds.append(e)
    # round 1: knights differ on line 2 of the template, so nothing of note is shared; round 2: a knight's prompt
    # extends its own round-1 prompt up to the end of the old transcript, and the page-aligned part of that is a hit
    for name, rnd, n_prompt, cached in seen:
        if rnd == 1:
            assert cached <= 80                                          # the chat wrapper and line 1 of the template
            continue
        a, b = prompts[(name, 1)], prompts[(name, 2)]
        common = next((i for i, (x, y) in enumerate(zip(a, b)) if x != y), min(len(a), len(b)))
        assert common > 3000                                            # template head + old transcript
        assert (common // 16) * 16 - 16 <= cached <= common
