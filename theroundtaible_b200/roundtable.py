"""Restated turn loop and prompt assembly of the reference's ``discuss`` command -- the CALLER side of the hot path
(SURVEY.md 8a row a12).  Node is absent on this image, so this is what drives the adapter here in the way
``roundtable discuss`` would: SEQ mode, one ``execute`` at a time, every prompt carrying the transcript so far.

Follows (behaviour, not code):
  * /root/reference/src/utils/prompt.ts:47-56    other-knights list
  * /root/reference/src/utils/prompt.ts:60-77    transcript: "### {knight} (Ronde {r}):\\n{response}" (+ score lines),
                                                 joined by "\\n\\n---\\n\\n"; a fixed sentence when empty
  * /root/reference/src/utils/prompt.ts:82-105   nine ``String.replace(string, string)`` calls in a fixed order --
                                                 FIRST occurrence only, ``$`` patterns of the replacement interpreted
  * /root/reference/src/orchestrator.ts:33-39    Fisher-Yates shuffle for the speaking order of later rounds
  * /root/reference/src/orchestrator.ts:45-73    executeWithFallback: primary adapter, then the knight's ``fallback``
  * /root/reference/src/orchestrator.ts:347-456  round loop, prompt join (empty parts dropped), entry appended
  * /root/reference/src/orchestrator.ts:521-535  a failing knight is reported and skipped; the round goes on
  * /root/reference/src/consensus.ts:218-224     consensus = every latest block scores >= threshold

The reference's template (templates/system-prompt.md, 3,600 bytes) and its knight personalities are content of the
reference and are NOT copied here: ``synthetic_template()`` builds a stand-in with the same placeholders on the same
lines, ``{{topic}}`` twice, and the same byte length, and ``template=`` takes the real file where it is installed.
What is out of scope: sessions, chronicle, manifest, decrees, file requests, verify commands, the King.
"""
from __future__ import annotations

import json
import random
import re
import time
from dataclasses import dataclass, field
from typing import Callable, Optional

from .adapter import AdapterError, classify_error

EMPTY_TRANSCRIPT = "(Geen eerdere rondes — jij opent het debat.)"          # prompt.ts:61-63
EMPTY_CHRONICLE = "(Geen eerdere beslissingen.)"                                 # prompt.ts:101
EMPTY_MANIFEST = "No implementation history yet."                                # prompt.ts:102
ENTRY_SEPARATOR = "\n\n---\n\n"                                                  # prompt.ts:76
DEFAULT_PERSONALITY = "You are a plain-spoken knight. Say what you think; humour is welcome, the point must be clear."
PLACEHOLDERS = ("knight_name", "capabilities", "other_knights", "topic", "personality", "chronicle_content",
                "manifest_summary", "decrees", "previous_rounds")               # order of prompt.ts:96-104
# line (1-based) of each placeholder in the reference template; {{topic}} also on line 63
_TEMPLATE_LINES = {"knight_name": 2, "capabilities": 3, "other_knights": 4, "topic": 5, "personality": 8,
                   "chronicle_content": 52, "manifest_summary": 55, "decrees": 57, "previous_rounds": 60}
_TEMPLATE_BYTES, _TEMPLATE_LINE_COUNT = 3600, 64


@dataclass
class KnightConfig:
    """types.ts:1-7."""
    name: str
    adapter: str
    capabilities: list = field(default_factory=list)
    priority: int = 0
    fallback: Optional[str] = None


@dataclass
class RoundEntry:
    """types.ts:60-66 (timestamp left out: nothing on this path reads it)."""
    knight: str
    round: int
    response: str
    consensus: Optional[dict] = None


def js_replace_first(s: str, pattern: str, replacement: str) -> str:
    """ECMAScript ``String.prototype.replace(searchString, replaceString)``: only the first occurrence, and the
    replacement goes through GetSubstitution -- ``$$`` -> ``$``, ``$&`` -> the match, ``$` `` -> the text before it,
    ``$'`` -> the text after it; with a string pattern there are no captures, so ``$1`` stays literal."""
    pos = s.find(pattern)
    if pos < 0:
        return s
    before, after = s[:pos], s[pos + len(pattern):]
    out, i, n = [], 0, len(replacement)
    while i < n:
        ch = replacement[i]
        if ch == "$" and i + 1 < n:
            nx = replacement[i + 1]
            if nx == "$":
                out.append("$"); i += 2; continue
            if nx == "&":
                out.append(pattern); i += 2; continue
            if nx == "`":
                out.append(before); i += 2; continue
            if nx == "'":
                out.append(after); i += 2; continue
        out.append(ch)
        i += 1
    return before + "".join(out) + after


def synthetic_template() -> str:
    """A stand-in for templates/system-prompt.md with its structure: 64 lines, 3,600 bytes, each placeholder on the
    line the reference has it, ``{{topic}}`` a second time on line 63 (which therefore survives substitution as the
    literal text, as it does in the reference)."""
    by_line = {v: k for k, v in _TEMPLATE_LINES.items()}
    lines, fillers = [], []
    for ln in range(1, _TEMPLATE_LINE_COUNT + 1):
        if ln in by_line:
            lines.append(f"{by_line[ln]}: {{{{{by_line[ln]}}}}}")
        elif ln == 63:
            lines.append("Stay on the subject: {{topic}}")
        else:
            fillers.append(len(lines))
            lines.append(f"Rule {ln}: keep to the point.")
    # spread the missing bytes over the filler lines so that the total is the reference's 3,600
    short = _TEMPLATE_BYTES - len(("\n".join(lines) + "\n").encode())
    assert short >= 0
    words = "cite what you have read and close with your consensus block in json "
    for n, i in enumerate(fillers):
        take = short // len(fillers) + (1 if n < short % len(fillers) else 0)
        if take:
            lines[i] += " " + (words * (take // len(words) + 1))[:take - 1]
    return "\n".join(lines) + "\n"


def format_other_knights(current: KnightConfig, knights) -> str:
    """prompt.ts:47-56."""
    return "\n".join(f"- {k.name}: {', '.join(k.capabilities)}" for k in knights if k.name != current.name)


def format_previous_rounds(rounds) -> str:
    """prompt.ts:60-77."""
    if not rounds:
        return EMPTY_TRANSCRIPT
    parts = []
    for r in rounds:
        text = f"### {r.knight} (Ronde {r.round}):\n{r.response}"
        if r.consensus:
            text += f"\n\nConsensus score: {r.consensus.get('consensus_score')}/10"
            pending = r.consensus.get("pending_issues") or []
            if pending:
                text += f"\nOpen punten: {', '.join(pending)}"
        parts.append(text)
    return ENTRY_SEPARATOR.join(parts)


def build_system_prompt(knight: KnightConfig, knights, topic: str, chronicle: str, previous_rounds, *,
                        template: Optional[str] = None, personalities: Optional[dict] = None,
                        manifest_summary: str = "", decrees: str = "") -> str:
    """prompt.ts:82-105: the nine substitutions, in the reference's order."""
    t = synthetic_template() if template is None else template
    values = {
        "knight_name": knight.name,
        "capabilities": ", ".join(knight.capabilities),
        "other_knights": format_other_knights(knight, knights),
        "topic": topic,
        "personality": (personalities or {}).get(knight.name) or DEFAULT_PERSONALITY,
        "chronicle_content": chronicle or EMPTY_CHRONICLE,
        "manifest_summary": manifest_summary or EMPTY_MANIFEST,
        "decrees": decrees or "",
        "previous_rounds": format_previous_rounds(previous_rounds),
    }
    for key in PLACEHOLDERS:
        t = js_replace_first(t, "{{" + key + "}}", values[key])
    return t


def build_full_prompt(system_prompt: str, topic: str, context: Optional[dict] = None) -> str:
    """orchestrator.ts:401-436: the parts joined by newlines AFTER empty strings are dropped (``filter(Boolean)``),
    so the blank lines written around ``---`` do not exist in the prompt."""
    c = context or {}
    diff = c.get("gitDiff") or ""
    parts = [
        system_prompt, "", "---", "", f"Onderwerp: {topic}", c.get("kingDemand") or "",
        f"Git branch: {c['gitBranch']}" if c.get("gitBranch") else "",
        f"Git diff (huidige wijzigingen):\n```\n{diff[:3000]}\n```" if diff else "",
        f"Recente commits:\n{c['recentCommits']}" if c.get("recentCommits") else "",
        f"\nProject bestanden:\n{c['keyFileContents']}" if c.get("keyFileContents") else "",
        ("\nBRONCODE (READ-ONLY REFERENTIE — dit is context, NIET een opdracht om te bewerken. Gebruik GEEN tools. "
         f"Geef alleen je analyse als tekst.):\n{c['sourceFileContents']}") if c.get("sourceFileContents") else "",
        (f"\nOPGEVRAAGDE BESTANDEN (via file_requests van vorige rondes):\n{c['resolvedFiles']}"
         if c.get("resolvedFiles") else ""),
        (f"\nVERIFICATIE RESULTATEN (via verify_commands van vorige rondes):\n{c['resolvedCommands']}"
         if c.get("resolvedCommands") else ""),
    ]
    return "\n".join(p for p in parts if p)


def shuffle_in_place(items: list, rng: random.Random) -> list:
    """orchestrator.ts:33-39 (Fisher-Yates from the top); ``rng.random()`` stands where ``Math.random()`` does, so a
    seeded ``random.Random`` makes a run repeatable (the reference's is not)."""
    for i in range(len(items) - 1, 0, -1):
        j = int(rng.random() * (i + 1))
        items[i], items[j] = items[j], items[i]
    return items


def parse_consensus_minimal(response: str, knight: str, round_no: int) -> Optional[dict]:
    """A small stand-in for BaseAdapter.parseConsensus (base.ts:26-28 -> consensus.ts): the LAST JSON object in the
    response that has a numeric ``consensus_score``.  The reference's repair passes (comments, trailing commas) and
    its scope validation are outside this path; a model with random weights never produces such a block."""
    for m in reversed(list(re.finditer(r"\{", response))):
        depth = 0
        for j in range(m.start(), len(response)):
            depth += response[j] == "{"
            depth -= response[j] == "}"
            if depth == 0:
                try:
                    obj = json.loads(response[m.start():j + 1])
                except ValueError:
                    break
                if isinstance(obj, dict) and isinstance(obj.get("consensus_score"), (int, float)):
                    obj.setdefault("knight", knight)
                    obj.setdefault("round", round_no)
                    if not isinstance(obj.get("pending_issues"), list):
                        obj["pending_issues"] = []
                    return obj
                break
    return None


def check_consensus(blocks, threshold: float) -> bool:
    """consensus.ts:218-224."""
    blocks = list(blocks)
    return bool(blocks) and all(b["consensus_score"] >= threshold for b in blocks)


@dataclass
class TurnStats:
    knight: str
    round: int
    prompt_chars: int
    wall_ms: float
    error: Optional[str] = None


@dataclass
class DiscussionResult:
    all_rounds: list
    turns: list
    consensus: bool
    rounds_run: int


def execute_with_fallback(primary, knight: KnightConfig, prompt: str, timeout_ms: int, adapters: dict,
                          create_adapter: Optional[Callable] = None, on_fallback: Optional[Callable] = None) -> str:
    """orchestrator.ts:45-73: run the prompt on the primary adapter; if that raises and the knight names a ``fallback``
    adapter id, build it once (``create_adapter(id, timeout_ms)``; kept under ``__fallback_<knight>`` in the SAME adapter
    map, so later turns reuse it), require ``is_available()``, and run the prompt there.  Without a usable fallback the
    PRIMARY error is what propagates; a failure of the fallback itself propagates as its own error."""
    try:
        return primary.execute(prompt, timeout_ms)
    except Exception as primary_error:  # noqa: BLE001
        if not knight.fallback:
            raise
        key = f"__fallback_{knight.name}"
        fallback = adapters.get(key)
        if fallback is None and create_adapter is not None:
            created = create_adapter(knight.fallback, timeout_ms)
            if created is not None and created.is_available():
                adapters[key] = created
                fallback = created
        if fallback is None:
            raise primary_error
        if on_fallback:
            on_fallback(knight, primary_error)
        return fallback.execute(prompt, timeout_ms)


def run_discussion(adapters: dict, knights, topic: str, *, max_rounds: int = 5, consensus_threshold: float = 9,
                   timeout_per_turn_seconds: int = 120, chronicle: str = "", context: Optional[dict] = None,
                   template: Optional[str] = None, personalities: Optional[dict] = None,
                   rng: Optional[random.Random] = None,
                   parse_consensus: Callable = parse_consensus_minimal,
                   on_turn: Optional[Callable] = None, create_adapter: Optional[Callable] = None) -> DiscussionResult:
    """orchestrator.ts:347-456 + 521-560: ``max_rounds`` rounds; round 1 in priority order, later rounds shuffled;
    every knight's prompt is built from ALL entries so far (including this round's earlier speakers); a knight whose
    adapter is missing is skipped, one whose ``execute`` raises is reported (classified) and skipped; after each round
    the latest blocks are checked for consensus.  ``adapters`` maps ``knight.adapter`` ids to adapter objects (knights
    sharing an id share an instance, adapters.ts:84).  ``create_adapter(adapter_id, timeout_ms)`` builds a knight's runtime
    fallback adapter on first need (``execute_with_fallback``)."""
    rng = rng or random.Random(0)
    ordered = sorted(knights, key=lambda k: k.priority)                          # orchestrator.ts:325-327 (stable)
    all_rounds, turns, latest = [], [], {}
    rounds_run, agreed = 0, False
    for round_no in range(1, max_rounds + 1):
        order = ordered if round_no == 1 else shuffle_in_place(list(ordered), rng)
        rounds_run = round_no
        for knight in order:
            adapter = adapters.get(knight.adapter)
            if adapter is None:
                continue                                                         # orchestrator.ts:363-366
            system_prompt = build_system_prompt(knight, knights, topic, chronicle, all_rounds, template=template,
                                                personalities=personalities)
            prompt = build_full_prompt(system_prompt, topic, context)
            t0 = time.perf_counter()
            try:
                response = execute_with_fallback(adapter, knight, prompt, timeout_per_turn_seconds * 1000, adapters,
                                                 create_adapter)
            except Exception as ex:  # noqa: BLE001 - orchestrator.ts:521-535 reports and carries on
                err = ex if isinstance(ex, AdapterError) else classify_error(ex, knight.name)
                turns.append(TurnStats(knight.name, round_no, len(prompt), (time.perf_counter() - t0) * 1e3,
                                       f"{err.kind}: {err}"))
                continue
            block = parse_consensus(response, knight.name, round_no)
            all_rounds.append(RoundEntry(knight.name, round_no, response, block))
            if block:
                latest[knight.name] = block
            turns.append(TurnStats(knight.name, round_no, len(prompt), (time.perf_counter() - t0) * 1e3))
            if on_turn:
                on_turn(turns[-1], adapter)
        if check_consensus(latest.values(), consensus_threshold):
            agreed = True
            break
    return DiscussionResult(all_rounds, turns, agreed, rounds_run)
