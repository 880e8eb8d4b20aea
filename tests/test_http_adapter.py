"""The reference's LocalLlmAdapter restated as a Python HTTP client (theroundtaible_b200/http_adapter.py) against a
scripted server: request shapes, response field paths, budgets, retry and error texts of
/root/reference/src/adapters/local-llm.ts.  CPU only (the reference holds no test for this adapter at all)."""
import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

import pytest

from theroundtaible_b200 import AdapterError
from theroundtaible_b200.http_adapter import HttpLocalLlmAdapter


class Script:
    """What the fake server answers; records what it was asked."""

    def __init__(self):
        self.requests = []
        self.chat = lambda body: (200, {"choices": [{"message": {"content": "antwoord"}}]})
        self.ollama = lambda body: (200, {"message": {"content": "antwoord"}})
        self.show = lambda body: (200, {"model_info": {"general.name": "x", "llama.context_length": 32768}})
        self.models_status = 200


@pytest.fixture()
def server():
    sc = Script()

    class H(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def _send(self, code, obj):
            b = obj if isinstance(obj, bytes) else json.dumps(obj).encode()
            self.send_response(code)
            self.send_header("Content-Length", str(len(b)))
            self.end_headers()
            self.wfile.write(b)

        def do_GET(self):  # noqa: N802
            sc.requests.append(("GET", self.path, None))
            self._send(sc.models_status, {"data": []})

        def do_POST(self):  # noqa: N802
            body = json.loads(self.rfile.read(int(self.headers.get("Content-Length", "0"))) or b"{}")
            sc.requests.append(("POST", self.path, body))
            route = {"/v1/chat/completions": sc.chat, "/api/chat": sc.ollama, "/api/show": sc.show}[self.path]
            self._send(*route(body))

        def log_message(self, *a):
            pass

    srv = ThreadingHTTPServer(("127.0.0.1", 0), H)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    sc.url = f"http://127.0.0.1:{srv.server_address[1]}"
    yield sc
    srv.shutdown()


def test_openai_compat_request_carries_only_model_and_messages(server):
    ad = HttpLocalLlmAdapter(server.url + "///", "qwen", "Merlin")               # trailing slashes stripped (local-llm.ts:17)
    assert ad.is_available() and ad.execute("hallo") == "antwoord"
    method, path, body = server.requests[-1]
    assert (method, path) == ("POST", "/v1/chat/completions")
    assert body == {"model": "qwen", "messages": [{"role": "user", "content": "hallo"}]}    # no max_tokens / temperature / stream
    assert ad.get_max_source_chars() is None                                     # unknown source: the orchestrator's 200 kB cap
    assert HttpLocalLlmAdapter(server.url, "qwen", "M", "LM Studio").get_max_source_chars() == (16384 - 4096 - 3000) * 4


def test_ollama_route_detects_context_and_sizes_num_ctx(server):
    ad = HttpLocalLlmAdapter(server.url, "llama", "Gawain", "Ollama")
    assert ad.detect_context_window() == 32768                                   # first "*.context_length" number
    assert ad.get_max_source_chars() == (32768 - 7096) * 4
    assert ad.execute("x" * 1001) == "antwoord"
    _, path, body = server.requests[-1]
    assert path == "/api/chat" and body["stream"] is False
    assert body["options"] == {"num_ctx": 251 + 4096 + 512}                      # ceil(1001 / 4) + reserves
    server.show = lambda b: (200, {"model_info": {"llama.context_length": 4000}})
    ad.detect_context_window()
    assert ad.get_max_source_chars() == 2000 * 4                                 # the 2,000-token floor
    ad.execute("x" * 40000)
    assert server.requests[-1][2]["options"]["num_ctx"] == 4000                  # clamped to the detected window
    server.show = lambda b: (500, b"no")
    assert ad.detect_context_window() is None


def test_errors_are_classified_like_the_reference(server):
    ad = HttpLocalLlmAdapter(server.url, "qwen", "Percival", "LM Studio", timeout_ms=5000)
    server.chat = lambda b: (500, b"context length exceeded: prompt too long")
    with pytest.raises(AdapterError) as ei:
        ad.execute("y" * 400)
    assert "LM Studio context window too small (prompt needs ~100 tokens)" in str(ei.value) and ei.value.kind == "unknown"
    plain = HttpLocalLlmAdapter(server.url, "qwen", "Percival")
    with pytest.raises(AdapterError) as ei:
        plain.execute("y")
    assert "Local LLM error (500): context length exceeded" in str(ei.value) and ei.value.kind == "api"   # "500" -> api
    server.chat = lambda b: (200, {"choices": [{"message": {"content": ""}}]})
    with pytest.raises(AdapterError, match="Local LLM returned empty response"):
        plain.execute("y")
    server.chat = lambda b: (500, b"generation timed out after 120000 ms")
    with pytest.raises(AdapterError) as ei:
        plain.execute("y")
    assert ei.value.kind == "timeout"                                            # "timed out" is tested before "500"
    assert HttpLocalLlmAdapter("http://127.0.0.1:9", "m", "Kay").is_available() is False
    with pytest.raises(AdapterError):
        HttpLocalLlmAdapter("http://127.0.0.1:9", "m", "Kay").execute("y")


def test_one_retry_after_model_reloaded(server):
    calls = []

    def flaky(body):
        calls.append(1)
        return (500, b"Model reloaded.") if len(calls) == 1 else (200, {"choices": [{"message": {"content": "tweede poging"}}]})
    server.chat = flaky
    ad = HttpLocalLlmAdapter(server.url, "qwen", "Bors", reload_wait_s=0.01)
    assert ad.execute("y") == "tweede poging" and len(calls) == 2
    calls.clear()
    server.chat = lambda b: (calls.append(1), (500, b"Model reloaded."))[1]
    with pytest.raises(AdapterError):
        ad.execute("y")
    assert len(calls) == 2                                                       # exactly one retry
