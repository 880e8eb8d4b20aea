"""CPU oracle for the local-model generation path (TEST INFRASTRUCTURE ONLY).

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it, and there only as the checker or as the timed
CPU arm.  The product (``theroundtaible_b200``) never imports this package and
fails loudly when its CUDA library is missing.

What it restates
----------------
The reference (``/root/reference/src/adapters/local-llm.ts:112,159``) forwards
each knight prompt over HTTP to an external Ollama / LM Studio / OpenAI-compatible
server.  That server is NOT in the reference repository, not listed in
``package.json:40-46`` and not version-pinned, so the reference itself holds no
arithmetic and no golden vector for this path: **parity is unpinned by the
reference**.  The oracle therefore restates the published decoder-only
transformer algorithm those servers run (Llama-3.x / Qwen2.5 architectures) and is
pinned instead against Hugging Face ``transformers`` 5.5.0
(``LlamaForCausalLM`` / ``Qwen2ForCausalLM``) run in the build container:
``tests/golden/make_golden.py`` generated the committed fixtures in
``tests/golden/*.npz`` and ``tests/test_oracle_golden.py`` checks the oracle
against them.
"""
