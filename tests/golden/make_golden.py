"""Generate tests/golden/*.npz by running Hugging Face transformers (the published algorithm the
oracle restates) in the BUILD container.  Run once:  python tests/golden/make_golden.py

The reference (/root/reference) holds no golden vector for generation (its only test,
tests/manual-adapter-test.mjs:1-28, drives vendor CLIs), so these fixtures pin the oracle
against HF ``LlamaForCausalLM`` / ``Qwen2ForCausalLM`` fp32 on the SAME seeded weights.
Each fixture stores: arch name, seed, input ids, HF logits for every position (fp32), and the
HF inv_freq vector.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle.archs import ARCHS  # noqa: E402
from oracle.model import make_state_dict  # noqa: E402


def hf_model(a, sd):
    from transformers import LlamaConfig, LlamaForCausalLM, Qwen2Config, Qwen2ForCausalLM
    common = dict(vocab_size=a.vocab, hidden_size=a.hidden, intermediate_size=a.ffn, num_hidden_layers=a.layers,
                  num_attention_heads=a.n_q, num_key_value_heads=a.n_kv, head_dim=a.head_dim,
                  rms_norm_eps=a.rms_eps, tie_word_embeddings=a.tied, max_position_embeddings=4096,
                  attention_dropout=0.0)
    if a.rope_factor > 0:
        rope = dict(rope_type="llama3", rope_theta=a.rope_theta, factor=a.rope_factor,
                    low_freq_factor=a.rope_low_freq, high_freq_factor=a.rope_high_freq,
                    original_max_position_embeddings=a.rope_orig_ctx)
    else:
        rope = dict(rope_type="default", rope_theta=a.rope_theta)
    if a.family == "llama":
        cfg = LlamaConfig(**common, rope_parameters=rope, attention_bias=False, mlp_bias=False)
        m = LlamaForCausalLM(cfg)
    else:
        cfg = Qwen2Config(**common, rope_parameters=rope, use_sliding_window=False)
        m = Qwen2ForCausalLM(cfg)
    m = m.to(torch.float32).eval()
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x))
    hsd = {"model.embed_tokens.weight": t(sd["embed"]), "model.norm.weight": t(sd["final_norm"]),
           "lm_head.weight": t(sd["lm_head"])}
    names = {"q": "self_attn.q_proj.weight", "k": "self_attn.k_proj.weight", "v": "self_attn.v_proj.weight",
             "o": "self_attn.o_proj.weight", "gate": "mlp.gate_proj.weight", "up": "mlp.up_proj.weight",
             "down": "mlp.down_proj.weight", "in_norm": "input_layernorm.weight",
             "post_norm": "post_attention_layernorm.weight", "q_bias": "self_attn.q_proj.bias",
             "k_bias": "self_attn.k_proj.bias", "v_bias": "self_attn.v_proj.bias"}
    for l in range(a.layers):
        for k, hn in names.items():
            if f"l{l}.{k}" in sd:
                hsd[f"model.layers.{l}.{hn}"] = t(sd[f"l{l}.{k}"])
    missing, unexpected = m.load_state_dict(hsd, strict=False)
    assert not unexpected, unexpected
    assert all("rotary" in k or "inv_freq" in k for k in missing), missing
    return m


def main():
    out_dir = os.path.dirname(os.path.abspath(__file__))
    torch.manual_seed(0)
    for name, seed, T in [("tiny-llama", 7, 96), ("tiny-llama-hd128", 3, 80), ("tiny-qwen", 11, 96),
                          ("tiny-qwen-hd128", 5, 72)]:
        a = ARCHS[name]
        sd = make_state_dict(a, seed)
        m = hf_model(a, sd)
        rng = np.random.default_rng(1234 + seed)
        ids = rng.integers(0, a.vocab, size=T, dtype=np.int64)
        with torch.no_grad():
            logits = m(torch.from_numpy(ids)[None]).logits[0].float().numpy()
        inv = m.model.rotary_emb.inv_freq.float().numpy()
        np.savez_compressed(os.path.join(out_dir, f"hf_{name}.npz"), arch=name, seed=seed, ids=ids,
                            logits=logits.astype(np.float32), inv_freq=inv)
        print(name, logits.shape, float(np.abs(logits).max()))


if __name__ == "__main__":
    main()
