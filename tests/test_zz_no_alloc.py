"""The allocation-free attention entry point (rt_k_attention_noalloc) against the ordinary one.

Some hosts of this library cannot call cudaMalloc: a serving framework that already owns all of HBM through
its own pool, or a latency-critical decode loop that cannot take the device-wide synchronisation an
allocation implies.  For them the split-KV scratch lives inside the library.  This test drives the new
entry point through the C-ABI with device buffers it owns (torch is only the owner of that memory), twice
in a row with different shapes, and compares with rt_k_attention -- same tolerance as the other attention
tests.  The file name sorts last so that a `-x` run has already executed every other test."""
import numpy as np
import pytest

from oracle import weights as W
from theroundtaible_b200 import engine as E

pytestmark = pytest.mark.gpu
PAGE = 16


def _paginate(k, v, lens):
    """[T][ctx][n_kv][hd] -> pool [pages][n_kv][16][hd] + page table, the layout the engine uses."""
    T, ctx, n_kv, hd = k.shape
    pps = -(-ctx // PAGE)
    table = np.zeros((T, pps), np.int32)
    pages_k, pages_v = [], []
    for t in range(T):
        for p in range(-(-lens[t] // PAGE)):
            table[t, p] = len(pages_k)
            blk_k = np.zeros((n_kv, PAGE, hd), np.uint16)
            blk_v = np.zeros((n_kv, PAGE, hd), np.uint16)
            n = min(PAGE, lens[t] - p * PAGE)
            blk_k[:, :n] = k[t, p * PAGE:p * PAGE + n].transpose(1, 0, 2)
            blk_v[:, :n] = v[t, p * PAGE:p * PAGE + n].transpose(1, 0, 2)
            pages_k.append(blk_k)
            pages_v.append(blk_v)
    return np.stack(pages_k), np.stack(pages_v), table


def _run_noalloc(torch, qb, kb, vb, lens, chunk):
    T, n_q, hd = qb.shape
    n_kv = kb.shape[2]
    pk, pv, table = _paginate(kb, vb, lens)
    dev = torch.device("cuda", 0)
    as_dev = lambda a: torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a).to(dev)
    dq, dk, dv, dt = as_dev(qb), as_dev(pk), as_dev(pv), as_dev(table)
    dpos = torch.tensor([l - 1 for l in lens], dtype=torch.int32, device=dev)
    dout = torch.full((T, n_q, hd), float("nan"), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    E.k_attention_noalloc(dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), dt.data_ptr(), table.shape[1], dpos.data_ptr(),
                          dout.data_ptr(), T, max(lens), n_q, n_kv, hd, chunk, device=0,
                          stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return dout.cpu().numpy()


def test_noalloc_attention_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    torch.cuda.set_device(0)
    cases = [(128, 32, 8, [300, 17, 256, 129, 64], 64), (64, 14, 2, [33, 90, 1], 32)]   # second call: other shape, stale scratch
    for i, (hd, n_q, n_kv, lens, chunk) in enumerate(cases):
        rng = np.random.default_rng(100 + i)
        T, ctx = len(lens), max(lens)
        qb = W.f32_to_bf16_bits(rng.standard_normal((T, n_q, hd)).astype(np.float32))
        kb = W.f32_to_bf16_bits(rng.standard_normal((T, ctx, n_kv, hd)).astype(np.float32))
        vb = W.f32_to_bf16_bits(rng.standard_normal((T, ctx, n_kv, hd)).astype(np.float32))
        want = E.k_attention(qb, kb, vb, lens, shared=0, chunk=chunk)
        got = _run_noalloc(torch, qb, kb, vb, lens, chunk)
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= 4e-3 * np.abs(want).max()


def test_noalloc_rejects_sizes_beyond_its_scratch():
    import torch
    torch.cuda.set_device(0)
    d = torch.zeros(16, dtype=torch.int32, device="cuda")
    with pytest.raises(E.EngineError, match="exceed the built-in maximum"):
        # 32 queries x 64 heads x 4096 chunks x 128 dims of partials: far beyond 4 MiB; rejected before any launch
        E.k_attention_noalloc(d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr(), 4096, d.data_ptr(), d.data_ptr(),
                              32, 65536, 64, 8, 128, 16)
