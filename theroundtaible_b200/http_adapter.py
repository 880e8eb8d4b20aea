"""The reference's ``LocalLlmAdapter`` as an HTTP client, restated in Python (Node is absent on this image).

This is the CLIENT half of the zero-TypeScript-change route: what the unmodified adapter
(/root/reference/src/adapters/local-llm.ts) does on the wire, so that the shim (http_shim.py) can be driven here exactly
as the reference would drive it -- availability probe, Ollama context detection, the 4-chars-per-token source budget,
both request formats, the one retry, and the error texts the orchestrator classifies.  Same constructor arguments,
same method meaning, same error behaviour:

  isAvailable            local-llm.ts:24-36    GET /v1/models, 3 s, any failure -> False
  detectContextWindow    local-llm.ts:43-49    only for source "Ollama": POST /api/show (local-llm.ts:205-235, 5 s)
  getMaxSourceChars      local-llm.ts:58-70    max(ctx - 4096 - 3000, 2000) * 4; "LM Studio" assumes 16,384; else None
  execute                local-llm.ts:72-89    Ollama -> /api/chat, otherwise /v1/chat/completions; one retry 3 s after an
                                               error whose text contains "Model reloaded"
  executeOllama          local-llm.ts:95-144   {model, messages, stream:false, options:{num_ctx}}, num_ctx = ceil(chars/4) +
                                               4096 + 512 clamped to the detected window; reads message.content
  executeOpenAICompat    local-llm.ts:150-199  {model, messages} and nothing else; reads choices[0].message.content
  isContextWindowError   local-llm.ts:240-248  substring test on the error body ("LM Studio" gets an explanatory text)

Every failure leaves through ``classify_error`` (src/utils/errors.ts:86-126), as in the reference.
"""
from __future__ import annotations

import json
import math
import re
import time
import urllib.error
import urllib.request
from typing import Optional

from .adapter import BaseAdapter, classify_error

RESPONSE_RESERVE_TOKENS = 4096      # local-llm.ts:65,102
OVERHEAD_RESERVE_TOKENS = 3000      # local-llm.ts:66
CHARS_PER_TOKEN = 4                 # local-llm.ts:68,101
LMSTUDIO_ASSUMED_CONTEXT = 16384    # local-llm.ts:60-61


class HttpLocalLlmAdapter(BaseAdapter):
    def __init__(self, endpoint: str, model: str, name: str, source: Optional[str] = None, timeout_ms: int = 120_000,
                 reload_wait_s: float = 3.0):
        self.endpoint = re.sub(r"/+$", "", endpoint)
        self.model = model
        self.name = name
        self.source = source                      # "Ollama" | "LM Studio" | None
        self.default_timeout = timeout_ms
        self.detected_context_tokens: Optional[int] = None
        self._reload_wait_s = reload_wait_s       # the reference waits 3 s; tests shorten it

    # -- probes -------------------------------------------------------------------------------------
    def is_available(self) -> bool:
        try:
            with urllib.request.urlopen(f"{self.endpoint}/v1/models", timeout=3.0) as r:
                return 200 <= r.status < 300
        except Exception:  # noqa: BLE001 - the reference swallows everything here
            return False

    def detect_context_window(self) -> Optional[int]:
        if self.source == "Ollama":
            self.detected_context_tokens = self._detect_ollama_context()
        return self.detected_context_tokens

    def _detect_ollama_context(self) -> Optional[int]:
        try:
            status, body = self._post("/api/show", {"name": self.model}, 5.0)
            if not 200 <= status < 300:
                return None
            info = json.loads(body).get("model_info")
            if isinstance(info, dict):
                for key, value in info.items():
                    if key.endswith(".context_length") and isinstance(value, (int, float)) and not isinstance(value, bool):
                        return int(value)
            return None
        except Exception:  # noqa: BLE001
            return None

    def get_max_source_chars(self) -> Optional[int]:
        ctx = self.detected_context_tokens or (LMSTUDIO_ASSUMED_CONTEXT if self.source == "LM Studio" else None)
        if not ctx:
            return None
        return max(ctx - RESPONSE_RESERVE_TOKENS - OVERHEAD_RESERVE_TOKENS, 2000) * CHARS_PER_TOKEN

    # -- execute -------------------------------------------------------------------------------------
    def execute(self, prompt: str, timeout_ms: Optional[int] = None) -> str:
        run = self._execute_ollama if self.source == "Ollama" else self._execute_openai_compat
        try:
            return run(prompt, timeout_ms)
        except Exception as ex:  # noqa: BLE001
            if "Model reloaded" in str(ex):
                time.sleep(self._reload_wait_s)
                return run(prompt, timeout_ms)
            raise

    def _post(self, path: str, obj: dict, timeout_s: float):
        req = urllib.request.Request(self.endpoint + path, data=json.dumps(obj).encode(),
                                     headers={"Content-Type": "application/json"}, method="POST")
        try:
            with urllib.request.urlopen(req, timeout=timeout_s) as r:
                return r.status, r.read().decode("utf-8", "replace")
        except urllib.error.HTTPError as he:         # a non-2xx answer is a response, not a transport failure
            return he.code, he.read().decode("utf-8", "replace")

    def _execute_ollama(self, prompt: str, timeout_ms: Optional[int]) -> str:
        timeout = self.default_timeout if timeout_ms is None else timeout_ms
        num_ctx = math.ceil(len(prompt) / CHARS_PER_TOKEN) + RESPONSE_RESERVE_TOKENS + 512
        if self.detected_context_tokens:
            num_ctx = min(num_ctx, self.detected_context_tokens)
        try:
            status, body = self._post("/api/chat", {"model": self.model, "messages": [{"role": "user", "content": prompt}],
                                                    "stream": False, "options": {"num_ctx": num_ctx}}, timeout / 1000.0)
            if not 200 <= status < 300:
                raise RuntimeError(f"Ollama error ({status}): {body}")
            content = (json.loads(body).get("message") or {}).get("content")
            if not content:
                raise RuntimeError("Ollama returned empty response")
            return content
        except Exception as ex:  # noqa: BLE001
            raise classify_error(ex, self.name) from ex

    def _execute_openai_compat(self, prompt: str, timeout_ms: Optional[int]) -> str:
        timeout = self.default_timeout if timeout_ms is None else timeout_ms
        try:
            # no max_tokens, no temperature, no stream: the server decides (local-llm.ts:156-165)
            status, body = self._post("/v1/chat/completions", {"model": self.model, "messages": [{"role": "user", "content": prompt}]},
                                      timeout / 1000.0)
            if not 200 <= status < 300:
                if self.source == "LM Studio" and self.is_context_window_error(body):
                    est = math.ceil(len(prompt) / CHARS_PER_TOKEN)
                    raise RuntimeError(f"LM Studio context window too small (prompt needs ~{est} tokens).\n"
                                       "  Fix: In LM Studio → Developer → Model Settings → increase Context Length.\n"
                                       "  Also uncheck the Response Limit, or set it higher.\n"
                                       "  Note: higher context = more VRAM. Find the sweet spot for your GPU.")
                raise RuntimeError(f"Local LLM error ({status}): {body}")
            choices = json.loads(body).get("choices") or []
            content = ((choices[0] if choices else {}).get("message") or {}).get("content")
            if not content:
                raise RuntimeError("Local LLM returned empty response")
            return content
        except Exception as ex:  # noqa: BLE001
            raise classify_error(ex, self.name) from ex

    @staticmethod
    def is_context_window_error(error_body: str) -> bool:
        low = error_body.lower()
        return (("n_keep" in low and "n_ctx" in low) or "context length exceeded" in low or "maximum context length" in low
                or "too many tokens" in low)
