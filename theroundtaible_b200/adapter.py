"""Host-side mirror of the reference's adapter interface for the local-llm path.

The reference is TypeScript and Node is absent on this image, so the host side above the C-ABI is
Python with the SAME names, argument meaning and error behaviour as:

  * ``BaseAdapter``       /root/reference/src/adapters/base.ts:10-29
  * ``LocalLlmAdapter``   /root/reference/src/adapters/local-llm.ts:6-249
  * ``createAdapter``     /root/reference/src/utils/adapters.ts:15-54 (the ``local-llm*`` branch)
  * ``classifyError``     /root/reference/src/utils/errors.ts:86-126

``B200LocalLlmAdapter.execute(prompt, timeout_ms)`` returns the generated text or raises
``AdapterError`` exactly where the reference adapter would reject its Promise.  The TypeScript
class a maintainer would drop into src/adapters/ (same shape, calling the N-API addon) is shown
in INTEGRATION.md.
"""
from __future__ import annotations

import os
import re
from typing import Optional

from .engine import RT_FLAG_PREFIX_CACHE, Engine, EngineError


class AdapterError(Exception):
    """errors.ts:57-80 -- AdapterError(adapter, message, {kind, hint, cause})."""

    def __init__(self, adapter: str, message: str, kind: str = "unknown", hint: Optional[str] = None, cause=None):
        super().__init__(message)
        self.adapter = adapter
        self.kind = kind
        self.hint = hint
        self.__cause__ = cause


def classify_error(error, adapter: str) -> AdapterError:
    """errors.ts:86-126: classify by lower-cased substring, in the reference's order."""
    if isinstance(error, AdapterError):
        return error
    msg = str(error)
    lower = msg.lower()
    cause = error if isinstance(error, BaseException) else None
    if any(s in lower for s in ("enoent", "not found", "not recognized", "command not found")):
        return AdapterError(adapter, f"{adapter} CLI not found: {msg}", "not_installed",
                            f"Is {adapter} installed and available in your PATH?", cause)
    if any(s in lower for s in ("timeout", "timed out", "etimedout")):
        return AdapterError(adapter, f"{adapter} timed out: {msg}", "timeout",
                            "Try increasing timeout_per_turn_seconds in .roundtable/config.json", cause)
    if any(s in lower for s in ("api key", "apikey", "unauthorized", "401", "403")):
        return AdapterError(adapter, f"{adapter} authentication failed: {msg}", "auth",
                            f"Check your API key or subscription for {adapter}.", cause)
    if any(s in lower for s in ("api error", "rate limit", "429", "too many requests", "500", "502", "503")):
        return AdapterError(adapter, f"{adapter} API error: {msg}", "api",
                            f"Wait a moment and try again, or check your {adapter} usage limits.", cause)
    return AdapterError(adapter, f"{adapter} error: {msg}", "unknown", None, cause)


class BaseAdapter:
    """base.ts:10-29."""

    name: str = ""

    def execute(self, prompt: str, timeout_ms: Optional[int] = None) -> str:  # pragma: no cover - abstract
        raise NotImplementedError

    def is_available(self) -> bool:  # pragma: no cover - abstract
        raise NotImplementedError

    def get_max_source_chars(self) -> Optional[int]:
        return None


class B200LocalLlmAdapter(BaseAdapter):
    """Drop-in for ``LocalLlmAdapter`` (local-llm.ts:6-22): same constructor arguments, with
    ``endpoint`` of the form ``b200://<device>`` selecting the in-process engine instead of an HTTP
    server.  ``source`` is accepted and ignored (there is one wire format: the C-ABI)."""

    def __init__(self, endpoint: str, model: str, name: str, source: Optional[str] = None, timeout_ms: int = 120_000,
                 *, seed: int = 0, max_batch: int = 8, max_ctx: int = 16384, max_new_tokens: int = 512,
                 ignore_eos: bool = False, engine: Optional[Engine] = None, temperature: float = 0.0, top_k: int = 0,
                 top_p: float = 0.0, sampling_seed: int = 0):
        self.endpoint = re.sub(r"/+$", "", endpoint)          # local-llm.ts:17 strips trailing slashes
        self.model = model
        self.name = name
        self.source = source
        self.default_timeout = timeout_ms
        self.max_new_tokens = max_new_tokens
        self.ignore_eos = ignore_eos
        # the reference request carries no sampling field (local-llm.ts:162-165); greedy unless the host opts in
        self._sampling = dict(temperature=temperature, top_k=top_k, top_p=top_p, seed=sampling_seed)
        self._engine = engine
        # the adapter keeps each knight's prompt KV between turns: its next prompt extends the last one (prompt.ts:60-77)
        self._engine_args = dict(seed=seed, max_batch=max_batch, max_ctx=max_ctx, flags=RT_FLAG_PREFIX_CACHE)
        self._load_error: Optional[BaseException] = None
        self.last_generation = None          # engine.Generation of the last execute(): token counts, timings, cache hits
        self.last_round = None               # the Generations of the last execute_round()

    # -- engine life cycle ----------------------------------------------------------------------
    def _device(self) -> int:
        m = re.match(r"^b200://(\d+)$", self.endpoint)
        if not m:
            raise EngineError(-1, f"endpoint '{self.endpoint}' not found: expected b200://<device ordinal>")
        return int(m.group(1))

    def _get_engine(self) -> Engine:
        if self._engine is None:
            if os.path.isdir(self.model):    # `model` names a Hugging Face checkpoint directory instead of an architecture
                self._engine = Engine.from_pretrained(self.model, device=self._device(), **self._engine_args)
            else:
                self._engine = Engine(self.model, device=self._device(), **self._engine_args)
        return self._engine

    # -- BaseAdapter ----------------------------------------------------------------------------
    def is_available(self) -> bool:
        """local-llm.ts:24-36: true when the backend answers; never throws."""
        try:
            self._get_engine()
            return True
        except Exception as ex:  # noqa: BLE001 - the reference swallows every error here
            self._load_error = ex
            return False

    def detect_context_window(self) -> Optional[int]:
        """local-llm.ts:43-49 (detectContextWindow): the engine knows its own context length."""
        try:
            return int(self._get_engine().info.context_length)
        except Exception:  # noqa: BLE001
            return None

    def get_max_source_chars(self) -> Optional[int]:
        """local-llm.ts:58-70: max(ctx - 4096 - 3000, 2000) * 4 chars.  NOTE: this backend's tokenizer is
        byte-level (1 token per UTF-8 byte), so the reference's 4-chars-per-token heuristic over-estimates
        what fits; the budget therefore uses 1 char per token."""
        ctx = self.detect_context_window()
        if not ctx:
            return None
        return max(ctx - min(4096, self.max_new_tokens) - 3000, 2000)

    def execute(self, prompt: str, timeout_ms: Optional[int] = None) -> str:
        """local-llm.ts:72-89 + 150-199: returns the completion text; raises AdapterError (classified)."""
        timeout = self.default_timeout if timeout_ms is None else timeout_ms
        try:
            gen = self._get_engine().generate(prompt, max_new_tokens=self.max_new_tokens, ignore_eos=self.ignore_eos,
                                              deadline_ms=int(timeout), **self._sampling)
            self.last_generation = gen
            if not gen.text:
                raise RuntimeError("Local LLM returned empty response")      # local-llm.ts:189-191
            return gen.text
        except Exception as ex:  # noqa: BLE001
            raise classify_error(ex, self.name) from ex

    # -- north-star capability (no reference counterpart today) --------------
    def execute_round(self, prompts, timeout_ms: Optional[int] = None):
        """All knights of one blind round in ONE batched generation; the common token prefix is prefilled
        once.  Returns one text per prompt."""
        timeout = self.default_timeout if timeout_ms is None else timeout_ms
        try:
            gens = self._get_engine().generate_batch(list(prompts), max_new_tokens=self.max_new_tokens,
                                                     ignore_eos=self.ignore_eos, deadline_ms=int(timeout), **self._sampling)
            self.last_round = gens
            return [g.text for g in gens]
        except Exception as ex:  # noqa: BLE001
            raise classify_error(ex, self.name) from ex


def create_adapter(adapter_id: str, config: dict, timeout_ms: Optional[int] = None, **engine_kw):
    """adapters.ts:45-54: any id starting with ``local-llm`` whose adapter_config entry has endpoint +
    model builds the local adapter; anything else is not this backend's business (returns None)."""
    entry = (config.get("adapter_config") or {}).get(adapter_id)
    if adapter_id.startswith("local-llm") and entry and "endpoint" in entry and "model" in entry:
        if str(entry["endpoint"]).startswith("b200://"):
            return B200LocalLlmAdapter(entry["endpoint"], entry["model"], entry.get("name") or adapter_id,
                                       entry.get("source"), timeout_ms if timeout_ms is not None else 120_000,
                                       **engine_kw)
    return None
