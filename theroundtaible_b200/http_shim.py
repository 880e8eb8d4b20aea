"""OpenAI-compatible HTTP shim over the native engine: the route that leaves every TypeScript file untouched.

The UNMODIFIED reference adapter (``LocalLlmAdapter``, /root/reference/src/adapters/local-llm.ts) becomes a client of
the B200 engine when its ``adapter_config`` entry points here:

  GET  /v1/models              availability probe (local-llm.ts:24-36) and ``init`` auto-detection
                               (src/utils/local-detect.ts:43-71): 200 {"data":[{"id": model}]}
  POST /v1/chat/completions    {model, messages:[{role:"user",content}]} (local-llm.ts:159-165) ->
                               200 {"choices":[{"message":{"role":"assistant","content": text}}]} (local-llm.ts:184-188)
  POST /api/show               {"name": model} -> {"model_info": {"b200.context_length": N}} (local-llm.ts:205-235),
                               for entries that declare ``"source": "Ollama"``.  N is an EFFECTIVE window: the adapter
                               budgets its source text at 4 chars per token -- (N - 4096 - 3000) * 4 chars
                               (local-llm.ts:58-70) -- but this backend's tokenizer is byte-level, 1 token per char, so N
                               is chosen such that the budget, the template and the transcript still fit the real window
                               (``effective_context``)
  POST /api/chat               Ollama-native form of the same request (local-llm.ts:112-133): {"message":{"content"}}

Failures answer 500 with the engine's error text as body, which the adapter wraps into
``Local LLM error (500): <text>`` (local-llm.ts:181).  Run:  python -m theroundtaible_b200.http_shim --model llama-3.2-1b
"""
from __future__ import annotations

import argparse
import json
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

from .engine import RT_FLAG_PREFIX_CACHE, Engine, EngineError

# what the reference adapter assumes when `source` is "LM Studio" (local-llm.ts:60-61): a 16,384-token window, i.e. a
# source budget of (16384 - 4096 - 3000) * 4 = 37,152 chars; with no `source` at all it assumes nothing and the
# orchestrator's own 200,000-char cap applies (orchestrator.ts:283)
LMSTUDIO_SOURCE_CHARS = (16384 - 4096 - 3000) * 4
PROMPT_OVERHEAD_TOKENS = 3000 * 4   # the adapter's own allowance for template, chronicle and rounds: 3,000 tokens at 4 chars each


def effective_context(max_ctx: int, max_new_tokens: int) -> int:
    """The context length /api/show reports.  The unmodified adapter turns a window of N tokens into a source budget of
    max(N - 4096 - 3000, 2000) * 4 CHARS (local-llm.ts:58-70).  Here a char is a token, so N must satisfy
    (N - 7096) * 4 + overhead + max_new_tokens <= max_ctx.  Never below the adapter's own floor (2,000 * 4 chars needs
    N >= 9,096 to be exceeded; smaller windows simply report the floor's equivalent)."""
    budget_chars = max(max_ctx - max_new_tokens - PROMPT_OVERHEAD_TOKENS, 0)
    return 7096 + budget_chars // 4


def make_server(engine: Engine, model: str, host: str = "127.0.0.1", port: int = 11435, max_new_tokens: int = 512,
                ignore_eos: bool = False, request_timeout_s: float = 120.0) -> ThreadingHTTPServer:
    """``request_timeout_s`` is the engine deadline of every request (default = the reference's
    ``timeout_per_turn_seconds`` of 120, src/commands/init.ts:212): the adapter aborts its fetch after that long
    (local-llm.ts:151-153), and a generation nobody is waiting for must not hold the engine."""
    lock = threading.Lock()     # one engine, one call at a time (the reference never overlaps execute() calls)

    def complete(body: dict) -> str:
        msgs = body.get("messages") or []
        if not msgs or not isinstance(msgs[-1], dict) or "content" not in msgs[-1]:
            raise ValueError("messages[-1].content missing")
        # the reference adapter sends none of these (local-llm.ts:162-165) -> greedy, `max_new_tokens`; other OpenAI /
        # Ollama clients may: honour the fields this backend has a kernel for
        opts = body.get("options") if isinstance(body.get("options"), dict) else body
        kw = dict(max_new_tokens=int(body.get("max_tokens") or opts.get("num_predict") or max_new_tokens),
                  temperature=float(opts.get("temperature") or 0.0), top_k=int(opts.get("top_k") or 0),
                  top_p=float(opts.get("top_p") or 0.0), seed=int(opts.get("seed") or 0))
        deadline = time.monotonic() + request_timeout_s
        # a request that waited for the engine past its own deadline is dropped instead of run for a client that has gone
        if not lock.acquire(timeout=max(request_timeout_s, 0.001)):
            raise EngineError(-4, f"generation timed out after {int(request_timeout_s * 1000)} ms waiting for the engine")
        try:
            left_ms = int((deadline - time.monotonic()) * 1000)
            if left_ms <= 0:
                raise EngineError(-4, f"generation timed out after {int(request_timeout_s * 1000)} ms waiting for the engine")
            gen = engine.generate(str(msgs[-1]["content"]), ignore_eos=ignore_eos, deadline_ms=left_ms, **kw)
            stats["requests"] += 1
            stats["prompt_tokens"] += gen.n_prompt_tokens
            stats["prefix_cached_tokens"] += gen.prefix_cached_tokens
            return gen.text
        finally:
            lock.release()

    stats = {"requests": 0, "prompt_tokens": 0, "prefix_cached_tokens": 0}

    class Handler(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def _send(self, code: int, obj):
            b = obj if isinstance(obj, bytes) else json.dumps(obj).encode()
            self.send_response(code)
            self.send_header("Content-Type", "application/json" if not isinstance(obj, bytes) else "text/plain")
            self.send_header("Content-Length", str(len(b)))
            self.end_headers()
            self.wfile.write(b)

        def do_GET(self):  # noqa: N802
            if self.path.rstrip("/") == "/v1/models":
                self._send(200, {"object": "list", "data": [{"id": model, "object": "model", "owned_by": "roundtable-b200"}]})
            elif self.path.rstrip("/") == "/stats":     # not part of the reference's wire format: cache hit rate for operators
                self._send(200, dict(stats, hit_rate=stats["prefix_cached_tokens"] / max(1, stats["prompt_tokens"])))
            else:
                self._send(404, b"not found")

        def do_POST(self):  # noqa: N802
            try:
                n = int(self.headers.get("Content-Length", "0"))
                body = json.loads(self.rfile.read(n) or b"{}")
                path = self.path.rstrip("/")
                if path == "/v1/chat/completions":
                    text = complete(body)
                    self._send(200, {"object": "chat.completion", "model": model,
                                     "choices": [{"index": 0, "message": {"role": "assistant", "content": text},
                                                  "finish_reason": "stop"}]})
                elif path == "/api/chat":
                    self._send(200, {"model": model, "message": {"role": "assistant", "content": complete(body)}, "done": True})
                elif path == "/api/show":
                    self._send(200, {"model_info": {"b200.context_length": effective_context(int(engine.info.context_length),
                                                                                             max_new_tokens)}})
                else:
                    self._send(404, b"not found")
            except Exception as ex:  # noqa: BLE001 - the text is what the adapter classifies (errors.ts:86-126)
                self._send(500, str(ex).encode("utf-8", "replace"))

        def log_message(self, *a):  # noqa: D102
            pass

    srv = ThreadingHTTPServer((host, port), Handler)
    srv.stats = stats
    return srv


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3.2-1b")
    ap.add_argument("--checkpoint", default=None, help="Hugging Face checkpoint directory (config.json + *.safetensors); "
                    "without it the model has seeded random weights")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--port", type=int, default=11435)
    # default window: the "LM Studio" route's fixed 37,152-char source budget + 12,000 chars of prompt overhead + the reply,
    # in BYTE tokens (50,688), rounded up
    ap.add_argument("--max-ctx", type=int, default=51200,
                    help="context window in byte-level tokens (1 per UTF-8 byte).  The unmodified adapter sends up to 37,152 chars "
                         "of source with source='LM Studio', (N-7096)*4 chars with source='Ollama' (N = what /api/show reports, "
                         "scaled down here to fit), and up to 200,000 chars with no source: size this for the route in use")
    ap.add_argument("--max-new-tokens", type=int, default=512)
    ap.add_argument("--request-timeout", type=float, default=120.0, help="seconds per request (reference default per turn: 120)")
    ap.add_argument("--knights", type=int, default=8, help="prompts kept in the cross-round prefix cache next to the running one")
    args = ap.parse_args()
    if args.checkpoint:
        raise SystemExit("--checkpoint: the built-in tokenizer is byte-level and does not match a checkpoint's vocabulary; "
                         "checkpoints are served through rt_generate_ids (theroundtaible_b200.engine), not this text route")
    # the cross-round prefix cache (a knight's next prompt extends its last one, prompt.ts:60-77) needs room for every
    # knight's last prompt next to the running sequence
    from .engine import arch_config
    a = arch_config(args.model)
    kv_tok = 2 * a.layers * a.n_kv * a.head_dim * 2
    eng = Engine(args.model, device=args.device, max_batch=1, max_ctx=args.max_ctx, flags=RT_FLAG_PREFIX_CACHE,
                 kv_pool_bytes=(args.knights + 1) * args.max_ctx * kv_tok)
    srv = make_server(eng, args.model, port=args.port, max_new_tokens=args.max_new_tokens, request_timeout_s=args.request_timeout)
    print(f"roundtable-b200 serving {args.model} on http://127.0.0.1:{args.port}", flush=True)
    srv.serve_forever()


if __name__ == "__main__":
    main()
