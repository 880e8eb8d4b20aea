"""Tensor-parallel decode probe under torchrun: every rank builds its shard, runs one BATCH round and prints the step time;
with RT_LIB pointing at the trace build, rank 0 dumps the in-kernel timeline to gpurun_out/<tag>_trace_rank0.bin
(scripts/trace_chain.py reads it).   usage: torchrun ... tp_probe.py <model> <knights> <shared> <own> <new> [tag]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
arg = lambda i, d: type(d)(sys.argv[i]) if len(sys.argv) > i else d
model, knights, shared, own, new, tag = arg(1, "llama-3.1-8b"), arg(2, 5), arg(3, 8192), arg(4, 128), arg(5, 24), arg(6, "tp")
os.environ.setdefault("NCCL_DEBUG", "WARN")
if rank == 0 and os.environ.get("RT_LIB"):
    os.environ["RT_TRACE_FILE"] = f"gpurun_out/{tag}_trace_rank0.bin"
from theroundtaible_b200 import engine as E
torch.cuda.set_device(local)
kw = {}
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(E.nccl_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    kw = dict(tp_rank=rank, tp_size=world, nccl_id=bytes(idt.cpu().numpy().tobytes()))
eng = E.Engine(model, seed=0, device=local, max_batch=knights, max_ctx=shared + own + new + 64, **kw)
rng = np.random.default_rng(7)
common = rng.integers(0, eng.vocab, shared).tolist()
prompts = [common + rng.integers(0, eng.vocab, own).tolist() for _ in range(knights)]
eng.generate_ids(prompts, max_new_tokens=4)            # warm-up (graph capture)
g = eng.generate_ids(prompts, max_new_tokens=new)
if rank == 0:
    print(f"{model} tp{world} knights {knights} ctx {shared}+{own}: prefill_ms {g[0].prefill_ms:.1f} us/step "
          f"{1e3 * g[0].decode_ms / max(1, g[0].decode_steps):.1f} weights/step {eng.info.weight_bytes_per_step / 1e9:.2f} GB", flush=True)
eng.close()
if world > 1:
    dist.destroy_process_group()
