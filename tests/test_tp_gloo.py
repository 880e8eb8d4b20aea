"""world_size-2 gloo test (CPU) of the tensor-parallel PARTITIONING the engine uses: the sharded
oracle forward with all-reduce must equal the unsharded oracle, and the vocab-parallel arg-max must equal argmax."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle.archs import ARCHS
from oracle.model import KVState, OracleModel
from oracle.tp import argmax_tp, forward_tp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, name, ids, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    m = OracleModel(ARCHS[name], 5, max_pos=128)
    kv = KVState()
    a = forward_tp(m, ids[:20], kv, rank, world)
    b = forward_tp(m, ids[20:], kv, rank, world)          # incremental continuation on the sharded KV
    local = torch.cat([a, b])
    toks = argmax_tp(local, rank, world)
    q.put((rank, local.numpy(), toks))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-qwen"])
def test_sharded_forward_equals_unsharded(name):
    a = ARCHS[name]
    ids = np.random.default_rng(3).integers(0, a.vocab, 33).tolist()
    ref = OracleModel(a, 5, max_pos=128).forward(ids, KVState(), round_kv=True, round_act=True).numpy()
    ctx = mp.get_context("spawn")
    q, port = ctx.Queue(), _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, ids, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict()
    for _ in range(2):
        r, logits, toks = q.get(timeout=120)
        got[r] = (logits, toks)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = np.concatenate([got[0][0], got[1][0]], axis=-1)
    assert np.abs(full - ref).max() <= 2e-3 * np.abs(ref).max()     # fp32 reduction order + bf16 re-rounding points
    assert got[0][1] == got[1][1]                                   # every rank picks the same tokens
    agree = np.mean(np.array(got[0][1]) == ref.argmax(-1))
    assert agree >= 0.9
