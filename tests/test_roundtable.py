"""CPU tests of the restated turn loop / prompt assembly (theroundtaible_b200/roundtable.py; reference:
src/orchestrator.ts:347-456, src/utils/prompt.ts:47-105).  Adapters are fakes: no engine, no GPU."""
from __future__ import annotations

import random

import pytest

from theroundtaible_b200 import roundtable as R
from theroundtaible_b200.adapter import AdapterError


class FakeAdapter:
    def __init__(self, name, reply=None, fail_on=()):
        self.name, self.prompts, self.reply, self.fail_on = name, [], reply, set(fail_on)

    def execute(self, prompt, timeout_ms=None):
        self.prompts.append((prompt, timeout_ms))
        if len(self.prompts) in self.fail_on:
            raise RuntimeError("request timed out after 120000ms")
        return self.reply(len(self.prompts)) if self.reply else f"answer {len(self.prompts)} of {self.name}"


KNIGHTS = [R.KnightConfig("Gawain", "local-llm-b", ["review"], 2), R.KnightConfig("Merlin", "local-llm-a", ["code", "docs"], 1),
           R.KnightConfig("Percival", "local-llm-c", ["tests"], 3)]


def test_js_replace_first_follows_ecmascript():
    assert R.js_replace_first("a {{x}} b {{x}}", "{{x}}", "V") == "a V b {{x}}"            # first occurrence only
    assert R.js_replace_first("a {{x}} b", "{{x}}", "$$") == "a $ b"
    assert R.js_replace_first("a {{x}} b", "{{x}}", "<$&>") == "a <{{x}}> b"
    assert R.js_replace_first("a {{x}} b", "{{x}}", "$`|$'") == "a a | b b"
    assert R.js_replace_first("a {{x}} b", "{{x}}", "$1 $0 $") == "a $1 $0 $ b"            # no captures: literal
    assert R.js_replace_first("nothing here", "{{x}}", "V") == "nothing here"


def test_synthetic_template_has_the_reference_structure():
    t = R.synthetic_template()
    assert len(t.encode()) == 3600 and t.count("\n") == 64                                # 3,582 chars / 3,600 bytes
    lines = t.split("\n")
    for key, ln in R._TEMPLATE_LINES.items():
        assert "{{" + key + "}}" in lines[ln - 1]
    assert t.count("{{topic}}") == 2 and "{{topic}}" in lines[62]


def test_system_prompt_substitution_order_and_second_topic_survives():
    p = R.build_system_prompt(KNIGHTS[1], KNIGHTS, "Use $& in Makefiles?", "", [])
    assert "knight_name: Merlin" in p and "capabilities: code, docs" in p
    assert "- Gawain: review\n- Percival: tests" in p and "- Merlin" not in p
    assert "topic: Use {{topic}} in Makefiles?" in p          # `$&` in the topic re-inserts the matched placeholder ...
    assert p.count("{{topic}}") == 2                          # ... next to the untouched second occurrence on line 63
    assert R.EMPTY_TRANSCRIPT in p and R.EMPTY_CHRONICLE in p and R.EMPTY_MANIFEST in p
    assert "{{previous_rounds}}" not in p and "{{decrees}}" not in p
    plain = R.build_system_prompt(KNIGHTS[1], KNIGHTS, "caching", "", [])
    assert plain.count("{{topic}}") == 1 and "Stay on the subject: {{topic}}" in plain     # second occurrence stays literal


def test_transcript_format():
    rounds = [R.RoundEntry("Merlin", 1, "first"), R.RoundEntry("Gawain", 1, "second", {"consensus_score": 7, "pending_issues": ["a", "b"]}),
              R.RoundEntry("Percival", 2, "third", {"consensus_score": 9, "pending_issues": []})]
    assert R.format_previous_rounds(rounds) == (
        "### Merlin (Ronde 1):\nfirst\n\n---\n\n### Gawain (Ronde 1):\nsecond\n\nConsensus score: 7/10\nOpen punten: a, b"
        "\n\n---\n\n### Percival (Ronde 2):\nthird\n\nConsensus score: 9/10")


def test_full_prompt_drops_empty_parts():
    assert R.build_full_prompt("SYS", "t") == "SYS\n---\nOnderwerp: t"
    full = R.build_full_prompt("SYS", "t", {"gitBranch": "main", "gitDiff": "x" * 4000, "recentCommits": "c1"})
    assert full.startswith("SYS\n---\nOnderwerp: t\nGit branch: main\nGit diff (huidige wijzigingen):\n```\n" + "x" * 3000 + "\n```\n")
    assert full.endswith("Recente commits:\nc1")


def test_turn_loop_order_transcript_growth_and_timeouts():
    adapters = {k.adapter: FakeAdapter(k.name) for k in KNIGHTS}
    res = R.run_discussion(adapters, KNIGHTS, "topic", max_rounds=3, timeout_per_turn_seconds=77, rng=random.Random(5))
    assert [e.knight for e in res.all_rounds[:3]] == ["Merlin", "Gawain", "Percival"]       # round 1: by priority
    assert sorted(e.knight for e in res.all_rounds[3:6]) == ["Gawain", "Merlin", "Percival"]
    assert [e.round for e in res.all_rounds] == [1, 1, 1, 2, 2, 2, 3, 3, 3]
    assert res.rounds_run == 3 and not res.consensus and len(res.turns) == 9
    # the t-th turn of the session sees exactly t earlier entries, same-round speakers included
    seen = {k.adapter: iter(adapters[k.adapter].prompts) for k in KNIGHTS}
    by_name = {k.name: k.adapter for k in KNIGHTS}
    for t, e in enumerate(res.all_rounds):
        prompt, timeout_ms = next(seen[by_name[e.knight]])
        assert timeout_ms == 77_000
        assert prompt.count("### ") == t and (R.EMPTY_TRANSCRIPT in prompt) == (t == 0)
        assert prompt.endswith("\n---\nOnderwerp: topic")
    # a knight's next prompt extends its previous one up to the end of the old transcript (what the prefix cache reuses)
    p0, p1 = adapters["local-llm-a"].prompts[0][0], adapters["local-llm-a"].prompts[1][0]
    head = p0[:p0.index(R.EMPTY_TRANSCRIPT)]
    assert p1.startswith(head)


def test_failing_knight_is_classified_and_skipped_and_missing_adapter_ignored():
    adapters = {"local-llm-a": FakeAdapter("Merlin", fail_on={2}), "local-llm-b": FakeAdapter("Gawain")}   # Percival's is missing
    res = R.run_discussion(adapters, KNIGHTS, "topic", max_rounds=2, rng=random.Random(1))
    assert [e.knight for e in res.all_rounds].count("Merlin") == 1 and "Percival" not in [e.knight for e in res.all_rounds]
    failed = [t for t in res.turns if t.error]
    assert len(failed) == 1 and failed[0].knight == "Merlin" and failed[0].error.startswith("timeout:")


def test_consensus_stops_the_loop():
    block = 'I agree.\n```json\n{"consensus_score": 9, "agrees_with": ["all"], "pending_issues": []}\n```'
    adapters = {k.adapter: FakeAdapter(k.name, reply=lambda n: block if n >= 2 else "not yet {broken json") for k in KNIGHTS}
    res = R.run_discussion(adapters, KNIGHTS, "topic", max_rounds=5)
    assert res.consensus and res.rounds_run == 2
    assert res.all_rounds[-1].consensus["consensus_score"] == 9 and res.all_rounds[0].consensus is None
    assert "Consensus score: 9/10" in R.format_previous_rounds(res.all_rounds)
    assert not R.check_consensus([], 9) and not R.check_consensus([{"consensus_score": 8}], 9)


def test_shuffle_is_the_reference_fisher_yates():
    class Fixed(random.Random):
        def __init__(self, vals):
            super().__init__(0)
            self.vals = iter(vals)

        def random(self):
            return next(self.vals)
    # i = 2: j = floor(0.9 * 3) = 2 (no swap); i = 1: j = floor(0.1 * 2) = 0 (swap 0 and 1)
    assert R.shuffle_in_place(["a", "b", "c"], Fixed([0.9, 0.1])) == ["b", "a", "c"]
    assert R.shuffle_in_place(["a", "b", "c"], Fixed([0.0, 0.99])) == ["c", "b", "a"]


def test_execute_with_fallback_follows_the_reference():
    """orchestrator.ts:45-73: primary first; on failure the knight's fallback adapter is created once, must be available,
    is kept under __fallback_<knight>, and answers this and later turns; without one the PRIMARY error propagates."""
    class Fake:
        def __init__(self, reply=None, fail=None, available=True):
            self.reply, self.fail, self.available, self.calls = reply, fail, available, 0

        def execute(self, prompt, timeout_ms=None):
            self.calls += 1
            if self.fail:
                raise RuntimeError(self.fail)
            return self.reply

        def is_available(self):
            return self.available

    primary = Fake(fail="usage limit reached (429)")
    spare = Fake(reply="van de reserve")
    made = []

    def create(adapter_id, timeout_ms):
        made.append(adapter_id)
        return spare
    k = R.KnightConfig("Merlin", "local-llm-a", ["x"], 0, fallback="local-llm-b")
    adapters = {"local-llm-a": primary}
    assert R.execute_with_fallback(primary, k, "p", 1000, adapters, create) == "van de reserve"
    assert made == ["local-llm-b"] and adapters["__fallback_Merlin"] is spare
    assert R.execute_with_fallback(primary, k, "p", 1000, adapters, create) == "van de reserve"
    assert made == ["local-llm-b"] and spare.calls == 2                          # created once, reused
    # no fallback configured / fallback unavailable / no factory: the primary error is what the caller sees
    plain = R.KnightConfig("Kay", "local-llm-a", ["x"], 1)
    with pytest.raises(RuntimeError, match="usage limit"):
        R.execute_with_fallback(primary, plain, "p", 1000, {}, create)
    k2 = R.KnightConfig("Bors", "local-llm-a", ["x"], 2, fallback="local-llm-c")
    with pytest.raises(RuntimeError, match="usage limit"):
        R.execute_with_fallback(primary, k2, "p", 1000, {}, lambda i, t: Fake(reply="x", available=False))
    with pytest.raises(RuntimeError, match="usage limit"):
        R.execute_with_fallback(primary, k2, "p", 1000, {}, None)
    # a failing fallback raises ITS error
    with pytest.raises(RuntimeError, match="ook stuk"):
        R.execute_with_fallback(primary, k2, "p", 1000, {}, lambda i, t: Fake(fail="ook stuk"))
    # and the turn loop uses it: the knight with a broken primary still speaks every round
    ok = Fake(reply="prima")
    knights = [R.KnightConfig("Merlin", "a", ["x"], 0, fallback="b"), R.KnightConfig("Kay", "c", ["x"], 1)]
    res = R.run_discussion({"a": Fake(fail="kapot"), "c": ok}, knights, "onderwerp", max_rounds=2,
                           create_adapter=lambda i, t: Fake(reply="reserve spreekt"))
    assert [t.error for t in res.turns] == [None] * 4
    assert [e.response for e in res.all_rounds if e.knight == "Merlin"] == ["reserve spreekt"] * 2
