"""Tensor parallel on real GPUs (needs >= 2 B200 on the box: `gpurun --gpus 2`; skipped otherwise).
One process per GPU, NCCL all-reduce per row-parallel projection, vocab-parallel arg-max.  TP=2 must reproduce
the oracle (and hence TP=1) under the same margin-aware greedy rule as the single-GPU parity tests."""
import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from oracle.archs import ARCHS
from oracle.model import KVState, OracleModel
from theroundtaible_b200 import engine as E

pytestmark = pytest.mark.gpu


def _rank_main(rank, tp, nccl_id, name, seed, prompts, n_new, mode, q, env=None):
    try:
        import os
        os.environ.update(env or {})
        torch.cuda.set_device(rank)
        with E.Engine(name, seed=seed, device=rank, max_batch=len(prompts), max_ctx=256, tp_rank=rank, tp_size=tp,
                      nccl_id=nccl_id) as eng:
            gens = eng.generate_ids(prompts, max_new_tokens=n_new)
            eng.seq_reset()
            logits = eng.step_logits(prompts[0], mode=mode)
        q.put((rank, [g.tokens for g in gens], logits, None))
    except Exception as ex:  # noqa: BLE001
        q.put((rank, None, None, repr(ex)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
# tiny-qwen has 7 q heads x 32 dims per rank: not a multiple of 64, so its prompts take the slab path (mode 0)
# third case: the NCCL all-reduce path instead of the fused peer-memory exchange (RT_TP_FUSED=0); fourth: the pull form of the
# fused exchange for decode steps too (RT_TP_PUSH=0; by default steps of <= 8 slots take the push form, larger ones the pull form)
@pytest.mark.parametrize("name,mode,env", [("tiny-llama", 1, None), ("tiny-qwen", 0, None), ("tiny-llama", 1, {"RT_TP_FUSED": "0"}),
                                           ("tiny-qwen", 0, {"RT_TP_PUSH": "0"})])
def test_tp2_generation_and_logits_match_oracle(name, mode, env):
    a = ARCHS[name]
    rng = np.random.default_rng(11)
    common = rng.integers(0, a.vocab, 48).tolist()
    prompts = [common + rng.integers(0, a.vocab, 9 + i).tolist() for i in range(3)]
    nccl_id = E.nccl_unique_id()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, nccl_id, name, 4, prompts, 20, mode, q, env)) for r in range(2)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(2):
        r, toks, logits, err = q.get(timeout=300)
        assert err is None, err
        res[r] = (toks, logits)
    for p in procs:
        p.join(timeout=60)
    assert res[0][0] == res[1][0]                                # ranks agree on every token
    oracle = OracleModel(a, 4, max_pos=256)
    ref = oracle.forward(prompts[0], KVState(), round_act=True).numpy()
    full = np.concatenate([res[0][1], res[1][1]], axis=-1)        # vocab-parallel shards
    assert (np.abs(full - ref).max(-1) / np.abs(ref).max(-1)).max() <= 1e-2
    from tests.test_gpu_parity import _greedy_compare
    for p, t in zip(prompts, res[0][0]):
        _greedy_compare(t, oracle, p, 20, err_bound=2.5e-2)


def _ckpt_rank_main(rank, tp, nccl_id, path, ids, q):
    try:
        torch.cuda.set_device(rank)
        kw = dict(tp_rank=rank, tp_size=tp, nccl_id=nccl_id) if tp > 1 else {}
        with E.Engine.from_pretrained(path, device=rank, max_batch=1, max_ctx=256, **kw) as eng:
            q.put((rank, eng.step_logits(ids, mode=1), None))
    except Exception as ex:  # noqa: BLE001
        q.put((rank, None, repr(ex)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_tp2_checkpoint_shards_match_hf(tmp_path):
    """A safetensors checkpoint loaded at TP=2: every rank takes its rows / columns of each tensor from the file, and
    the concatenated vocab shards equal the Hugging Face forward."""
    from tests.test_gpu_safetensors import _hf_model
    m = _hf_model("qwen2", tied=False, head_dim=64, seed=21)
    m.to(torch.bfloat16).save_pretrained(tmp_path, safe_serialization=True)
    m.to(torch.float32)
    ids = np.random.default_rng(5).integers(0, 512, 64).tolist()
    with torch.no_grad():
        want = m(torch.tensor(ids)[None]).logits[0].float().numpy()
    nccl_id = E.nccl_unique_id()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ckpt_rank_main, args=(r, 2, nccl_id, str(tmp_path), ids, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(2):
        r, logits, err = q.get(timeout=300)
        assert err is None, err
        res[r] = logits
    for p in procs:
        p.join(timeout=60)
    full = np.concatenate([res[0], res[1]], axis=-1)
    assert np.abs(full - want).max() <= 1e-2 * np.abs(want).max()
