"""The zero-TypeScript-change route: an HTTP client that speaks exactly what the unmodified reference adapter speaks
(/root/reference/src/adapters/local-llm.ts:24-36,150-199,205-235) against the shim over the native engine."""
import json
import threading
import urllib.error
import urllib.request

import pytest

from theroundtaible_b200 import engine as E
from theroundtaible_b200 import roundtable as R
from theroundtaible_b200.http_adapter import HttpLocalLlmAdapter
from theroundtaible_b200.http_shim import PROMPT_OVERHEAD_TOKENS, effective_context, make_server

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
        # the window reported is the EFFECTIVE one: the adapter budgets 4 chars per token, this tokenizer has 1 per byte
        assert any(k.endswith(".context_length") and v == effective_context(1024, 24) for k, v in data["model_info"].items())
        with pytest.raises(urllib.error.HTTPError) as ei:                             # prompt longer than the context
            _post(base + "/v1/chat/completions", {"model": "tiny-qwen", "messages": [{"role": "user", "content": "x" * 2000}]})
        assert ei.value.code == 500 and b"context length exceeded" in ei.value.read()
    finally:
        srv.shutdown()
        eng.close()


def _serve(eng, model, **kw):
    srv = make_server(eng, model, port=0, **kw)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    return srv, f"http://127.0.0.1:{srv.server_address[1]}"


def test_unmodified_adapter_budget_fits_the_byte_tokenizer():
    """The reference turns a window of N tokens into (N - 4096 - 3000) * 4 CHARS of source (local-llm.ts:58-70); with one
    token per byte that overflowed the real window (round-1 finding).  /api/show now reports an effective window, so a
    prompt carrying exactly the budget the adapter computes -- plus the template and the prompt scaffolding -- is served."""
    max_ctx, new = 28672, 32
    eng = E.Engine("tiny-qwen", seed=2, max_batch=1, max_ctx=max_ctx)
    srv, base = _serve(eng, "tiny-qwen", max_new_tokens=new)
    try:
        ad = HttpLocalLlmAdapter(base, "tiny-qwen", "Merlin", "Ollama")
        assert ad.is_available()
        n = ad.detect_context_window()
        assert n == effective_context(max_ctx, new) and n > 7096 + 2000
        budget = ad.get_max_source_chars()
        assert budget == (n - 7096) * 4 and budget + PROMPT_OVERHEAD_TOKENS + new <= max_ctx
        system = R.build_system_prompt(R.KnightConfig("Merlin", "local-llm-a", ["analysis"]), [], "cache voor de gateway?", "", [])
        prompt = R.build_full_prompt(system, "cache voor de gateway?", {"sourceFileContents": "s" * budget, "gitDiff": "d" * 3000})
        assert len(prompt) > budget + 3600 + 3000
        assert len(ad.execute(prompt)) > 0                                        # fits: 200, non-empty content
        with pytest.raises(Exception, match="context length exceeded"):          # four times that budget does not
            ad.execute(prompt + "s" * (3 * budget))
    finally:
        srv.shutdown()
        eng.close()


def test_discussion_through_the_shim_hits_the_prefix_cache_and_honours_the_deadline():
    """Five rounds of the restated turn loop, every turn an HTTP request in the reference's wire format: with the
    cross-round prefix cache on in the shim's engine a knight's round r+1 prompt reuses its round-r KV."""
    import random
    a = E.arch_config("tiny-llama")
    ctx = 16384
    eng = E.Engine("tiny-llama", seed=4, max_batch=1, max_ctx=ctx, flags=E.RT_FLAG_PREFIX_CACHE,
                   kv_pool_bytes=4 * ctx * 2 * a.layers * a.n_kv * a.head_dim * 2)
    srv, base = _serve(eng, "tiny-llama", max_new_tokens=24, ignore_eos=True)
    try:
        knights = [R.KnightConfig(n, f"local-llm-{i}", ["analysis"], i) for i, n in enumerate(["Merlin", "Gawain", "Kay"])]
        adapters = {k.adapter: HttpLocalLlmAdapter(base, "tiny-llama", k.name) for k in knights}
        res = R.run_discussion(adapters, knights, "moet de cache voor de gateway?", max_rounds=5, rng=random.Random(3))
        assert len(res.turns) == 15 and not any(t.error for t in res.turns)
        with urllib.request.urlopen(base + "/stats", timeout=10) as r:
            st = json.loads(r.read())
        assert st["requests"] == 15 and st["hit_rate"] > 0.5, st
    finally:
        srv.shutdown()
        eng.close()
    # a request whose deadline passes answers 500 with text the reference classifies as a timeout (errors.ts:102)
    eng = E.Engine("llama-3.2-1b", seed=0, max_batch=1, max_ctx=4096)
    srv, base = _serve(eng, "llama-3.2-1b", max_new_tokens=3000, ignore_eos=True, request_timeout_s=0.05)
    try:
        with pytest.raises(Exception) as ei:
            HttpLocalLlmAdapter(base, "llama-3.2-1b", "Lancelot").execute("x" * 200)
        assert ei.value.kind == "timeout" and "timed out" in str(ei.value)
    finally:
        srv.shutdown()
        eng.close()
