"""CPU reference server (TEST INFRASTRUCTURE; see oracle/__init__.py): the stand-in for the external
OpenAI-compatible server the reference adapter calls.

Wire format served (/root/reference/src/adapters/local-llm.ts:24-36,150-199):
  GET  /v1/models              -> 200 {"data":[{"id": model}]}
  POST /v1/chat/completions    {model, messages:[{role:"user",content}]} ->
                               {"choices":[{"message":{"role":"assistant","content": text}}]}
It is the arm `bench.py --impl reference` and `cpu_baseline` time on the host cores; it is never on the
product path.
"""
from __future__ import annotations

import json
import time
from http.server import BaseHTTPRequestHandler, HTTPServer

import torch

from . import tokenizer as T
from .archs import ARCHS
from .model import KVState, OracleModel, decode_step_batch


class OracleServer:
    def __init__(self, model: str, seed: int = 0, max_pos: int = 4096, max_new_tokens: int = 512, threads: int | None = None,
                 layers: int | None = None):
        """``layers``: keep only the first n decoder layers (models too big for host RAM: bench.py times a truncated
        depth at TRUE dimensions and extrapolates the layer loop, flagged as such)."""
        if threads:
            torch.set_num_threads(threads)
        self.model_name = model
        self.max_new_tokens = max_new_tokens
        arch = ARCHS[model]
        self.full_layers = arch.layers
        if layers and layers < arch.layers:
            arch = arch.with_layers(layers)
        self.m = OracleModel(arch, seed, max_pos=max_pos)

    # one request, greedy (what one knight turn costs on the CPU server)
    def complete_ids(self, ids, max_new_tokens=None):
        return self.m.generate(ids, max_new_tokens or self.max_new_tokens, round_act=True)

    def chat_completions(self, body: dict) -> dict:
        msgs = body.get("messages") or []
        if not msgs or "content" not in msgs[-1]:
            raise ValueError("messages[-1].content missing")
        ids = T.encode_chat(msgs[-1]["content"])
        out = self.complete_ids(ids)
        return {"object": "chat.completion", "model": self.model_name,
                "choices": [{"index": 0, "message": {"role": "assistant", "content": T.decode(out) or "<eos>"},
                             "finish_reason": "length"}]}

    # N concurrent requests with the shared prefix prefilled once and the decode projections batched:
    # the fairest CPU counterpart of the engine's BATCH mode.  Returns (ids per request, timings).
    def complete_batch_ids(self, prompts, max_new_tokens, shared: int = 0, prefill_chunk: int = 1024):
        t0 = time.perf_counter()
        self.m.layer_seconds = 0.0
        base = KVState()
        for s in range(0, shared, prefill_chunk):          # chunked: the score matrix of a chunk stays in cache-sized pieces
            self.m.forward(prompts[0][s:min(shared, s + prefill_chunk)], base, round_act=True, last_only=True)
        kvs, toks = [], []
        for p in prompts:
            kv = base.clone() if shared else KVState()
            lg = None
            for s in range(shared, len(p), prefill_chunk):
                lg = self.m.forward(p[s:s + prefill_chunk], kv, round_act=True, last_only=True)
            kvs.append(kv)
            toks.append(int(lg[-1].argmax()))
        t1 = time.perf_counter()
        prefill_layer_s = self.m.layer_seconds
        self.m.layer_seconds = 0.0
        outs = [[t] for t in toks]
        for _ in range(max_new_tokens - 1):
            lg = decode_step_batch(self.m, toks, kvs)
            toks = [int(x) for x in lg.argmax(-1)]
            for o, t in zip(outs, toks):
                o.append(t)
        t2 = time.perf_counter()
        return outs, {"prefill_s": t1 - t0, "decode_s": t2 - t1, "decode_steps": max_new_tokens - 1,
                      "prefill_layer_s": prefill_layer_s, "decode_layer_s": self.m.layer_seconds,
                      "layers": self.m.a.layers, "full_layers": self.full_layers}


def serve(model: str, port: int = 11435, seed: int = 0, max_new_tokens: int = 64):
    srv = OracleServer(model, seed, max_new_tokens=max_new_tokens)

    class H(BaseHTTPRequestHandler):
        def _send(self, code, obj):
            b = json.dumps(obj).encode()
            self.send_response(code)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(b)))
            self.end_headers()
            self.wfile.write(b)

        def do_GET(self):
            if self.path.rstrip("/") == "/v1/models":
                self._send(200, {"object": "list", "data": [{"id": srv.model_name, "object": "model"}]})
            else:
                self._send(404, {"error": "not found"})

        def do_POST(self):
            try:
                body = json.loads(self.rfile.read(int(self.headers.get("Content-Length", "0"))) or b"{}")
                if self.path.rstrip("/") == "/v1/chat/completions":
                    self._send(200, srv.chat_completions(body))
                else:
                    self._send(404, {"error": "not found"})
            except Exception as ex:  # noqa: BLE001
                self._send(500, {"error": str(ex)})

        def log_message(self, *a):
            pass

    HTTPServer(("127.0.0.1", port), H).serve_forever()
