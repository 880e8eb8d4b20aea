"""OpenAI-compatible HTTP shim over the native engine: the route that leaves every TypeScript file untouched.

The UNMODIFIED reference adapter (``LocalLlmAdapter``, /root/reference/src/adapters/local-llm.ts) becomes a client of
the B200 engine when its ``adapter_config`` entry points here:

  GET  /v1/models              availability probe (local-llm.ts:24-36) and ``init`` auto-detection
                               (src/utils/local-detect.ts:43-71): 200 {"data":[{"id": model}]}
  POST /v1/chat/completions    {model, messages:[{role:"user",content}]} (local-llm.ts:159-165) ->
                               200 {"choices":[{"message":{"role":"assistant","content": text}}]} (local-llm.ts:184-188)
  POST /api/show               {"name": model} -> {"model_info": {"b200.context_length": N}} (local-llm.ts:205-235),
                               for entries that declare ``"source": "Ollama"``
  POST /api/chat               Ollama-native form of the same request (local-llm.ts:112-133): {"message":{"content"}}

Failures answer 500 with the engine's error text as body, which the adapter wraps into
``Local LLM error (500): <text>`` (local-llm.ts:181).  Run:  python -m theroundtaible_b200.http_shim --model llama-3.2-1b
"""
from __future__ import annotations

import argparse
import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

from .engine import Engine


def make_server(engine: Engine, model: str, host: str = "127.0.0.1", port: int = 11435, max_new_tokens: int = 512,
                ignore_eos: bool = False) -> ThreadingHTTPServer:
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
        with lock:
            return engine.generate(str(msgs[-1]["content"]), ignore_eos=ignore_eos, **kw).text

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
                    self._send(200, {"model_info": {"b200.context_length": int(engine.info.context_length)}})
                else:
                    self._send(404, b"not found")
            except Exception as ex:  # noqa: BLE001 - the text is what the adapter classifies (errors.ts:86-126)
                self._send(500, str(ex).encode("utf-8", "replace"))

        def log_message(self, *a):  # noqa: D102
            pass

    return ThreadingHTTPServer((host, port), Handler)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3.2-1b")
    ap.add_argument("--checkpoint", default=None, help="Hugging Face checkpoint directory (config.json + *.safetensors); "
                    "without it the model has seeded random weights")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--port", type=int, default=11435)
    ap.add_argument("--max-ctx", type=int, default=16384)
    ap.add_argument("--max-new-tokens", type=int, default=512)
    args = ap.parse_args()
    if args.checkpoint:
        eng = Engine.from_pretrained(args.checkpoint, device=args.device, max_batch=8, max_ctx=args.max_ctx)
    else:
        eng = Engine(args.model, device=args.device, max_batch=8, max_ctx=args.max_ctx)
    srv = make_server(eng, args.model, port=args.port, max_new_tokens=args.max_new_tokens)
    print(f"roundtable-b200 serving {args.model} on http://127.0.0.1:{args.port}", flush=True)
    srv.serve_forever()


if __name__ == "__main__":
    main()
