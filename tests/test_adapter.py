"""Host-side mirror of the reference adapter interface (no GPU needed): error classification follows
src/utils/errors.ts:86-126, construction follows src/utils/adapters.ts:45-54 / local-llm.ts:15-22."""
import pytest
import torch

from theroundtaible_b200 import AdapterError, B200LocalLlmAdapter, classify_error, create_adapter

NO_GPU = not torch.cuda.is_available()


@pytest.mark.parametrize("msg,kind", [
    ("spawn claude ENOENT", "not_installed"), ("model 'x' not found", "not_installed"),
    ("generation timed out after 120001 ms (deadline 120000 ms)", "timeout"), ("ETIMEDOUT", "timeout"),
    ("Local LLM error (401): nope", "auth"), ("Local LLM error (503): busy", "api"), ("rate limit", "api"),
    ("This operation was aborted", "unknown"), ("context length exceeded: ...", "unknown"),
    ("CUDA error: out of memory", "unknown")])
def test_classify_error_matches_reference_order(msg, kind):
    err = classify_error(RuntimeError(msg), "Claude")
    assert isinstance(err, AdapterError) and err.kind == kind and err.adapter == "Claude"
    assert classify_error(err, "Other") is err                      # already classified -> returned as is


def test_classify_precedence_not_found_beats_timeout():
    # errors.ts checks not_installed first, then timeout, then auth, then api
    assert classify_error(RuntimeError("not found after timeout 401"), "k").kind == "not_installed"
    assert classify_error(RuntimeError("timeout 401 429"), "k").kind == "timeout"
    assert classify_error(RuntimeError("403 then 500"), "k").kind == "auth"


def test_create_adapter_registration_rule():
    cfg = {"adapter_config": {"local-llm-b200": {"endpoint": "b200://0///", "model": "llama-3.2-1b", "name": "Merlin"},
                              "local-llm-http": {"endpoint": "http://localhost:1234", "model": "x"},
                              "claude-cli": {"endpoint": "b200://0", "model": "x"}}}
    a = create_adapter("local-llm-b200", cfg, 5000)
    assert isinstance(a, B200LocalLlmAdapter) and a.name == "Merlin" and a.endpoint == "b200://0" and a.default_timeout == 5000
    assert create_adapter("local-llm-http", cfg) is None         # plain HTTP endpoints stay with the reference adapter
    assert create_adapter("claude-cli", cfg) is None
    assert create_adapter("local-llm-missing", cfg) is None
    b = create_adapter("local-llm-b200", {"adapter_config": {"local-llm-b200": {"endpoint": "b200://1", "model": "m"}}})
    assert b.name == "local-llm-b200" and b.default_timeout == 120_000


@pytest.mark.skipif(not NO_GPU, reason="checks the no-GPU failure mode")
def test_adapter_without_gpu_is_unavailable_and_execute_raises_classified():
    a = B200LocalLlmAdapter("b200://0", "tiny-llama", "Galahad")
    assert a.is_available() is False                              # local-llm.ts:24-36 never throws
    assert a.get_max_source_chars() is None
    with pytest.raises(AdapterError) as ei:
        a.execute("hallo")
    assert ei.value.adapter == "Galahad" and "no CUDA device" in str(ei.value)
    bad = B200LocalLlmAdapter("http://localhost:1234", "tiny-llama", "K")
    with pytest.raises(AdapterError) as ei:
        bad.execute("x")
    assert ei.value.kind == "not_installed"
