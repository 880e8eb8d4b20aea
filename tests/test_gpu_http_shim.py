"""The zero-TypeScript-change route: an HTTP client that speaks exactly what the unmodified reference adapter speaks
(/root/reference/src/adapters/local-llm.ts:24-36,150-199,205-235) against the shim over the native engine."""
import json
import threading
import urllib.error
import urllib.request

import pytest

from theroundtaible_b200 import engine as E
from theroundtaible_b200.http_shim import make_server

pytestmark = pytest.mark.gpu


def _post(url, obj):
    req = urllib.request.Request(url, data=json.dumps(obj).encode(), headers={"Content-Type": "application/json"}, method="POST")
    with urllib.request.urlopen(req, timeout=60) as r:
        return r.status, json.loads(r.read())


def test_reference_wire_format_round_trip():
    eng = E.Engine("tiny-qwen", seed=2, max_batch=1, max_ctx=1024)
    srv = make_server(eng, "tiny-qwen", port=0, max_new_tokens=24)
    port = srv.server_address[1]
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    try:
        base = f"http://127.0.0.1:{port}"
        with urllib.request.urlopen(base + "/v1/models", timeout=10) as r:          # isAvailable(): response.ok
            assert r.status == 200 and json.loads(r.read())["data"][0]["id"] == "tiny-qwen"
        prompt = "Je neemt deel aan een TheRoundtAIble discussie.\nJe naam is: Merlin"
        st, data = _post(base + "/v1/chat/completions", {"model": "tiny-qwen", "messages": [{"role": "user", "content": prompt}]})
        content = data["choices"][0]["message"]["content"]                            # local-llm.ts:188
        assert st == 200 and isinstance(content, str) and len(content) > 0
        assert content == eng.generate(prompt, max_new_tokens=24).text                # same engine, same greedy text
        st, data = _post(base + "/api/chat", {"model": "tiny-qwen", "messages": [{"role": "user", "content": prompt}], "stream": False,
                                              "options": {"num_ctx": 4096}})
        assert data["message"]["content"] == content                                  # local-llm.ts:133
        st, data = _post(base + "/api/show", {"name": "tiny-qwen"})
        assert any(k.endswith(".context_length") and v == 1024 for k, v in data["model_info"].items())
        with pytest.raises(urllib.error.HTTPError) as ei:                             # prompt longer than the context
            _post(base + "/v1/chat/completions", {"model": "tiny-qwen", "messages": [{"role": "user", "content": "x" * 2000}]})
        assert ei.value.code == 500 and b"context length exceeded" in ei.value.read()
    finally:
        srv.shutdown()
        eng.close()
