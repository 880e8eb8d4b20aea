#!/usr/bin/env python
"""Benchmark of the local-model generation path (BASELINE.json metric: decode tok/s + full-round latency,
3-8 knights, 1/2/4/8 B200, % of the HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload config3|...]

A STEP is one round of the roundtable in BATCH mode: the knight prompts of the round (shared prefix + per-knight
suffix, synthetic token ids) are prefilled and every knight generates `new_tokens` tokens greedily.

Headline workload = BASELINE.json configs[2] ("config3": 5 knights, Llama-3.1-8B, 8,192 shared + 128 own prompt
tokens, 512 new tokens), the largest configuration that fits one GPU.  `--gpus N` (torchrun, one rank per GPU) runs
the SAME model tensor-parallel over the N GPUs (strong scaling, `parallelism: tpN`): per layer two fused all-reduces
over NVLink peer memory, vocab-parallel sampling.  `--replicas` restores N independent copies (no collective).

  value        whole-job DECODE tok/s from the engine's CUDA events on its own stream (state resident in HBM)
  ms_per_step  full-round latency (prefill + 512 decode steps), barrier + synchronize on both sides, max over ranks
  e2e          the same round through the adapter API with HOST strings (tokenise -> H2D -> generate -> D2H -> text)
  extra        config 2 (3 knights, Llama-3.2-1B: last round's headline) and the 5-round SEQ discussion at N = 1;
               the BASELINE tensor-parallel configs where N matches (config 4 at N = 2, 4; config 5 at N = 8);
               tp_check = all ranks produce the same tokens, and the tensor-parallel logits of the 2-layer model at true
               dimensions sit within 1e-2 of the CPU oracle

--impl reference times the CPU stand-in of the external server the reference adapter calls (oracle/, `kind: port`)
on all host cores, on the SAME prompt shape; what does not fit the time / RAM budget is extrapolated and flagged
(decoder layers beyond the first `--cpu-layers`, decode steps beyond `--cpu-new-tokens`).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG", "WARN")      # no version banner on stdout: rank 0 prints exactly one JSON line

WORKLOADS = {
    # BASELINE.json configs[1]: 3 knights, Llama-3.2-1B bf16, shared system-prompt prefix, greedy
    "config2": dict(model="llama-3.2-1b", knights=3, shared=1024, own=64, new_tokens=512, max_ctx=2048, seed_idx=2),
    # BASELINE.json configs[2]: 5 knights, Llama-3.1-8B, 8k shared transcript prefix
    "config3": dict(model="llama-3.1-8b", knights=5, shared=8192, own=128, new_tokens=512, max_ctx=9216, seed_idx=3),
    # BASELINE.json configs[3]/[4]: tensor parallel only
    "config4": dict(model="qwen2.5-32b", knights=8, shared=2048, own=128, new_tokens=256, max_ctx=2560, seed_idx=4),
    "config5": dict(model="llama-3.1-70b", knights=8, shared=16384, own=128, new_tokens=256, max_ctx=16896, seed_idx=5),
    # smoke-sized
    "tiny": dict(model="tiny-llama", knights=3, shared=64, own=8, new_tokens=32, max_ctx=256, seed_idx=0),
}


def synth_prompts(w, vocab, rank=0):
    import numpy as np
    rng = np.random.default_rng(1234 + w["seed_idx"] + 1000 * rank)
    common = rng.integers(0, vocab, w["shared"]).tolist()
    return [common + rng.integers(0, vocab, w["own"]).tolist() for _ in range(w["knights"])]


def synth_text_prompts(w, rank=0):
    """Host strings of the same shape for the e2e arm (byte-level tokenizer: 1 token per ASCII char; the
    chat wrapper adds 2 leading tokens that join the shared prefix and 1 trailing token)."""
    import numpy as np
    rng = np.random.default_rng(4321 + w["seed_idx"] + 1000 * rank)
    alpha = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz ABCDEFGHIJKLMNOPQRSTUVWXYZ.,\n", dtype=np.uint8)
    common = bytes(alpha[rng.integers(0, len(alpha), w["shared"] - 2)]).decode()
    return [common + bytes(alpha[rng.integers(0, len(alpha), w["own"] - 1)]).decode() for _ in range(w["knights"])]


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:  # noqa: BLE001
                self.proc.kill()
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = max(mx, float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
        sm.sort()
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p["bf16_tflops"]), float(p["bf16_tflops_sustained"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, 1590.0, 1400.0, "fallback"


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


# =====================================================================================================
# CPU arm (the oracle port: the stand-in for the external server the reference delegates to)
# =====================================================================================================
def _host_ram_bytes():
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable"):
                    return int(line.split()[1]) * 1024
    except Exception:  # noqa: BLE001
        pass
    return 32 << 30


def cpu_round(w, cpu_layers, cpu_new_tokens, srv=None):
    """One round of workload `w` on the CPU port at the FULL prompt shape.  Bounded by (a) the first `cpu_layers` decoder
    layers at true dimensions, the layer loop scaled linearly to the full depth, and (b) `cpu_new_tokens` decode steps, the
    rest of the round extrapolated at the measured per-step time.  Returns (server, dict)."""
    import torch
    from oracle.archs import ARCHS
    from oracle.server import OracleServer
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    a = ARCHS[w["model"]]
    per_layer = 4 * (a.hidden * (a.n_q + 2 * a.n_kv) * a.head_dim + a.n_q * a.head_dim * a.hidden + 3 * a.hidden * a.ffn)
    fixed = 4 * a.vocab * a.hidden * (1 if a.tied else 2)
    layers = a.layers
    n_params = a.layers * per_layer // 4 + fixed // 4
    if cpu_layers == 0 and n_params > 3e9:          # auto: full depth for the small models, one layer for 8B and up
        cpu_layers = 1
    if cpu_layers and cpu_layers < layers:
        layers = cpu_layers
    while layers > 1 and fixed + layers * per_layer > 0.6 * _host_ram_bytes():       # fp32 copies of the bf16 weights
        layers //= 2
    new_tokens = min(w["new_tokens"], cpu_new_tokens)
    if srv is None:
        srv = OracleServer(w["model"], seed=0, max_pos=w["shared"] + w["own"] + new_tokens + 8, layers=layers)
    prompts = synth_prompts(w, srv.m.a.vocab)
    shared_pages = (w["shared"] // 16) * 16
    t0 = time.perf_counter()
    _, t = srv.complete_batch_ids(prompts, new_tokens, shared=shared_pages)
    wall = time.perf_counter() - t0
    scale = t["full_layers"] / t["layers"]
    # layer loop scaled to the full depth; embedding, final norm, lm_head and sampling are depth-independent
    prefill_s = (t["prefill_s"] - t["prefill_layer_s"]) + t["prefill_layer_s"] * scale
    step_s = ((t["decode_s"] - t["decode_layer_s"]) + t["decode_layer_s"] * scale) / max(t["decode_steps"], 1)
    full_steps = w["new_tokens"] - 1
    round_s = prefill_s + step_s * full_steps
    extrap = {}
    if t["layers"] < t["full_layers"]:
        extrap["layers"] = f"{t['layers']} of {t['full_layers']} decoder layers run at true dimensions, layer-loop time x {scale:g}"
    if new_tokens < w["new_tokens"]:
        extrap["new_tokens"] = f"{new_tokens} of {w['new_tokens']} tokens per knight decoded, the rest at the measured per-step time"
    desc = (f"{w['knights']} knights, {w['model']}, {w['shared']} shared + {w['own']} own prompt tokens (full prompt shape), "
            f"{new_tokens} new tokens per knight; oracle port, fp32 compute over bf16 weights, fp16 projection inputs, batched "
            f"decode, torch threads={cores}; measured prefill {t['prefill_s']:.2f}s decode {t['decode_s']:.2f}s"
            + ("; EXTRAPOLATED: " + "; ".join(extrap.values()) if extrap else ""))
    return srv, {"decode_tok_per_s": w["knights"] / step_s, "round_ms": 1e3 * round_s, "prefill_ms": 1e3 * prefill_s,
                 "e2e_tok_per_s": w["knights"] * w["new_tokens"] / round_s, "sample_wall_s": wall, "cores": cores,
                 "sample": desc, "extrapolated": extrap or None}


def run_reference(args, w):
    """--impl reference: K timed steps, each one bounded sample of the same round on the host cores (rank 0 only)."""
    rank, _, _ = dist_env()
    if rank != 0:
        return
    from oracle.build import build as obuild
    try:
        obuild()
    except Exception:  # noqa: BLE001
        pass
    srv, acc, last = None, [], None
    for i in range(args.warmup + args.steps):
        srv, r = cpu_round(w, args.cpu_layers, args.cpu_new_tokens, srv)
        if i >= args.warmup:
            acc.append(r)
        last = r
    n = len(acc)
    dec = n / sum(1.0 / r["decode_tok_per_s"] for r in acc)           # total tokens / total time
    round_ms = sum(r["round_ms"] for r in acc) / n
    e2e = w["knights"] * w["new_tokens"] / (round_ms * 1e-3)
    cpu = {"value": dec, "unit": "tok/s", "cores": last["cores"], "kind": "port", "sample": last["sample"],
           "extrapolated": last["extrapolated"], "sample_wall_s_per_step": sum(r["sample_wall_s"] for r in acc) / n}
    line = {"impl": "reference", "metric": "decode_tok_per_s", "value": dec, "unit": "tok/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round_ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.workload, w, w["model"], "host CPU cores (no GPU)"),
            "cpu_baseline": cpu,
            "e2e": {"value": e2e, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "note": "whole round (prefill + all decode steps) on the CPU port; see cpu_baseline.extrapolated"},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(name, w, model, parallelism):
    return {"workload": (f"{name}: {w['knights']} knights, {w['model']} bf16 weights, {w['shared']} shared + {w['own']} own "
                         f"prompt tokens, {w['new_tokens']} new tokens per knight, greedy, BATCH round"),
            "weights": "seeded random-init (Irwin-Hall approximation of N(0, 0.02)), bf16",
            "l2": "inputs larger than L2: one decode step streams the model's weights (>= 2.4 GB per GPU) against a 126 MB L2",
            "parallelism": parallelism, "model": model}


# =====================================================================================================
# GPU arm
# =====================================================================================================
class Dist:
    def __init__(self):
        import torch
        self.torch = torch
        self.rank, self.world, self.local = dist_env()
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (this backend has no CPU fallback)")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

    def barrier(self):
        if self.world > 1:
            self.torch.distributed.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = self.torch.tensor(values, dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.torch.distributed.all_reduce(t, op=self.torch.distributed.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def nccl_id(self, E):
        torch = self.torch
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if self.rank == 0:
            idt.copy_(torch.frombuffer(bytearray(E.nccl_unique_id()), dtype=torch.uint8))
        torch.distributed.broadcast(idt, 0)
        return bytes(idt.cpu().numpy().tobytes())

    def same_on_all_ranks(self, ints):
        torch = self.torch
        if self.world == 1:
            return True
        t = torch.tensor(ints, dtype=torch.int64, device="cuda")
        lo, hi = t.clone(), t.clone()
        torch.distributed.all_reduce(lo, op=torch.distributed.ReduceOp.MIN)
        torch.distributed.all_reduce(hi, op=torch.distributed.ReduceOp.MAX)
        return bool((lo == hi).all().item())

    def close(self):
        if self.world > 1:
            self.torch.distributed.destroy_process_group()


def make_engine(D, E, model, w, tp, **kw):
    if tp > 1:
        return E.Engine(model, seed=0, device=D.local, max_batch=w["knights"], max_ctx=w["max_ctx"], tp_rank=D.rank, tp_size=tp,
                        nccl_id=D.nccl_id(E), **kw)
    return E.Engine(model, seed=0, device=D.local, max_batch=w["knights"], max_ctx=w["max_ctx"], **kw)


def measure(D, E, w, model, tp, steps, warmup, *, e2e=True, seq=False, sampler=None):
    """Time `steps` BATCH rounds of workload `w` (after `warmup`), then the same through the adapter with host strings."""
    from theroundtaible_b200 import B200LocalLlmAdapter
    eng = make_engine(D, E, model, w, tp)
    replica_rank = 0 if tp > 1 else D.rank
    prompts = synth_prompts(w, eng.vocab, replica_rank)
    nt = w["new_tokens"]
    for _ in range(warmup):
        eng.generate_ids(prompts, max_new_tokens=nt, ignore_eos=True)
    if sampler:
        sampler.start()
    D.barrier()
    t0 = time.perf_counter()
    dec_ms = pre_ms = 0.0
    toks = launches = steps_dec = 0
    tokens_sig = []
    for _ in range(steps):
        gens = eng.generate_ids(prompts, max_new_tokens=nt, ignore_eos=True)
        dec_ms += gens[0].decode_ms
        pre_ms += gens[0].prefill_ms
        steps_dec += gens[0].decode_steps
        toks += sum(g.decode_steps for g in gens)
        launches += gens[0].gpu_launches
        tokens_sig = [t for g in gens for t in g.tokens]
    D.barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    out = {"tokens_same_on_all_ranks": D.same_on_all_ranks(tokens_sig) if tp > 1 else None}

    e2e_ms, e2e_tokens, h2d, d2h = 0.0, 0, 0, 0
    if e2e:
        texts = synth_text_prompts(w, replica_rank)
        adapter = B200LocalLlmAdapter(f"b200://{D.local}", w["model"], "bench", max_new_tokens=nt, ignore_eos=True, engine=eng)
        for _ in range(min(warmup, 2)):
            adapter.execute_round(texts)
        D.barrier()
        e0 = time.perf_counter()
        for _ in range(steps):
            outs = adapter.execute_round(texts)
            e2e_tokens += len(outs) * nt
        D.barrier()
        e2e_ms = (time.perf_counter() - e0) * 1e3
        n_tok = sum(len(t.encode()) + 3 for t in texts)
        h2d = n_tok * 4 * 3 + 16 * (n_tok // 16 + len(texts))          # token, sequence, position per prompt token + work items
        d2h = w["knights"] * nt * 4
    if sampler:
        out["clocks"] = sampler.stop()

    seq_res = None
    if seq:
        seq_dec_ms, seq_tok = 0.0, 0
        s0 = time.perf_counter()
        for p in prompts:
            g = eng.generate_ids([p], max_new_tokens=nt, ignore_eos=True)[0]
            seq_dec_ms += g.decode_ms
            seq_tok += g.decode_steps
        seq_res = {"decode_tok_per_s": seq_tok / (seq_dec_ms * 1e-3), "round_ms": (time.perf_counter() - s0) * 1e3,
                   "note": "SEQ = one knight per call (B=1), what orchestrator.ts:361-442 issues"}

    dec_ms_m, wall_ms_m, e2e_ms_m, pre_ms_m = D.max_over_ranks([dec_ms, wall_ms, e2e_ms, pre_ms])
    replicas = 1 if tp > 1 else D.world
    info, cfg = eng.info, eng.cfg
    hbm, tf_burst, tf_sust, peak_src = measured_peaks()
    # algorithmic bytes of one decode step ON ONE GPU: this rank's weights once + shared prefix KV once + own KV
    mean_own_ctx = w["own"] + nt / 2.0
    shared_pages = (w["shared"] // 16) * 16
    kv_b = info.kv_bytes_per_token * (shared_pages + w["knights"] * (w["shared"] - shared_pages + mean_own_ctx))
    bytes_step = info.weight_bytes_per_step + kv_b
    step_ms = dec_ms_m / max(steps_dec, 1)
    achieved = bytes_step / (step_ms * 1e-3) / 1e9
    # prefill: FLOPs per GPU = 2 * this rank's projection weights * tokens + causal attention (QK^T and PV) of its heads
    pf_tokens = shared_pages + w["knights"] * (w["shared"] - shared_pages + w["own"])
    per_layer_w = (cfg.hidden * (cfg.n_q + 2 * cfg.n_kv) * cfg.head_dim + cfg.n_q * cfg.head_dim * cfg.hidden + 3 * cfg.hidden * cfg.ffn) // tp
    attn_pairs = shared_pages * (shared_pages + 1) / 2 + w["knights"] * sum(range(shared_pages + 1, w["shared"] + w["own"] + 1))
    pf_flops = cfg.layers * (2.0 * per_layer_w * pf_tokens + 4.0 * (cfg.n_q // tp) * cfg.head_dim * attn_pairs)
    pf_ms = pre_ms_m / steps
    pf_tflops = pf_flops / (pf_ms * 1e-3) / 1e12
    out.update({
        "value": replicas * toks / (dec_ms_m * 1e-3), "ms_per_step": wall_ms_m / steps, "step_ms": step_ms,
        "prefill_ms": pf_ms, "launches": launches, "bytes_step": bytes_step, "achieved_gbs": achieved, "frac": achieved / hbm,
        "peak": hbm, "peak_src": peak_src, "weight_bytes_per_step": int(info.weight_bytes_per_step), "n_params": int(info.n_params),
        "e2e": ({"value": replicas * e2e_tokens / (e2e_ms_m * 1e-3), "unit": "tok/s", "h2d_bytes_per_step": h2d,
                 "d2h_bytes_per_step": d2h, "ms_per_round": e2e_ms_m / steps,
                 "note": "generated tokens / wall time of adapter.execute_round with host strings (tokenise, H2D, prefill, decode, D2H, detokenise)"}
                if e2e else None),
        "prefill": {"tokens_prefilled": pf_tokens, "ms": pf_ms, "tok_per_s": pf_tokens / (pf_ms * 1e-3), "tflops_per_gpu": pf_tflops,
                    "bound": "tensor", "frac_of_measured_burst": pf_tflops / tf_burst, "frac_of_measured_sustained": pf_tflops / tf_sust,
                    "note": "whole prefill phase (tcgen05 GEMMs + flash attention + norms" + (" + all-reduces" if tp > 1 else "") +
                            "), algorithmic FLOPs per GPU / CUDA-event time"},
        "seq": seq_res, "cfg": cfg,
    })
    eng.close()
    return out


def brief(m, w, model, tp):
    """Compact record of a secondary workload for `extra`."""
    return {"workload": workload_config("", w, model, f"tp{tp}" if tp > 1 else "single GPU")["workload"].lstrip(": "),
            "decode_tok_per_s": m["value"], "decode_ms_per_token_step": m["step_ms"], "round_ms": m["ms_per_step"],
            "prefill_ms": m["prefill_ms"], "prefill_tflops_per_gpu": m["prefill"]["tflops_per_gpu"],
            "roofline_frac_per_gpu": m["frac"], "achieved_gbs_per_gpu": m["achieved_gbs"],
            "e2e_tok_per_s": m["e2e"]["value"] if m["e2e"] else None, "seq": m["seq"],
            "tokens_same_on_all_ranks": m["tokens_same_on_all_ranks"]}


def tp_check(D, E, w, tp):
    """After the timed region: the headline architecture at TRUE dimensions and 2 layers (stated), tensor-parallel over the
    same ranks -- (a) every rank generates the same tokens, (b) the concatenated vocab shards of the teacher-forced logits,
    through the prefill kernels and through the decode kernels, sit within 1e-2 of max |logit| of the CPU oracle
    (the checker; rank 0 only).  The driver's GPU-test box has one GPU, so this is where the TP path is verified."""
    import numpy as np
    torch = D.torch
    model = w["model"] + "@L2"
    rng = np.random.default_rng(99)
    wt = dict(w, knights=3, max_ctx=512)
    eng = make_engine(D, E, model, wt, tp)
    vocab = eng.vocab
    ids = rng.integers(0, vocab, 96)
    prompts = [ids[:64].tolist() + rng.integers(0, vocab, 8 + i).tolist() for i in range(3)]
    gens = eng.generate_ids(prompts, max_new_tokens=16)
    same = D.same_on_all_ranks([t for g in gens for t in g.tokens])
    eng.seq_reset()
    l1 = eng.step_logits(ids, mode=1)
    eng.seq_reset()
    l0 = eng.step_logits(ids[:32], mode=0)
    eng.close()
    full = []
    for part in (l1, l0):
        t = torch.from_numpy(part).cuda()
        if tp > 1:
            parts = [torch.empty_like(t) for _ in range(tp)]
            torch.distributed.all_gather(parts, t)
            t = torch.cat(parts, dim=-1)
        full.append(t.cpu().numpy())
    res = None
    if D.rank == 0:
        from oracle.archs import ARCHS
        from oracle.model import KVState, OracleModel
        ref = OracleModel(ARCHS[w["model"]].with_layers(2), 0, max_pos=256).forward(ids, KVState(), round_act=True).numpy()
        sc = np.abs(ref).max(-1)
        e1 = float((np.abs(full[0] - ref).max(-1) / sc).max())
        e0 = float((np.abs(full[1] - ref[:32]).max(-1) / sc[:32]).max())
        res = {"model": model, "tp": tp, "tokens_same_on_all_ranks": same, "logits_rel_err_prefill_kernels": e1,
               "logits_rel_err_decode_kernels": e0, "tolerance": 1e-2, "ok": bool(same and e1 <= 1e-2 and e0 <= 1e-2),
               "note": "true dimensions, depth truncated to 2 layers (the CPU oracle cannot hold the full model in seconds)"}
    D.barrier()
    return res


def discuss_leg(w, model, device, rounds=5):
    """SEQ mode on the REAL prompt layout: the restated turn loop (theroundtaible_b200/roundtable.py, reference
    orchestrator.ts:347-456 + prompt.ts:60-105) drives one adapter per knight over ONE engine, `rounds` rounds, each
    turn's prompt = template + transcript so far.  Run twice: cross-call prefix cache on (the adapter's default) and
    off.  Reports wall time per round and the share of prompt tokens whose KV came from the cache."""
    import random
    from theroundtaible_b200 import B200LocalLlmAdapter
    from theroundtaible_b200 import engine as E
    from theroundtaible_b200 import roundtable as R
    names = ["Merlin", "Gawain", "Percival", "Lancelot", "Galahad", "Bors", "Kay", "Tristan"][:w["knights"]]
    knights = [R.KnightConfig(n, f"local-llm-{i}", ["analysis", "code review"], i) for i, n in enumerate(names)]
    nt = w["new_tokens"]
    out = {}
    for label, flags in (("prefix_cache_on", E.RT_FLAG_PREFIX_CACHE), ("prefix_cache_off", 0)):
        max_ctx = 3800 + rounds * len(knights) * (nt * 10 + 64) + nt      # a generated id detokenises to <= 10 bytes = 10 tokens
        a = E.arch_config(model)
        kv_tok = 2 * a.layers * a.n_kv * a.head_dim * 2
        # room for every knight's last prompt to stay cached next to the running sequence
        eng = E.Engine(model, seed=0, device=device, max_batch=2, max_ctx=max_ctx,
                       kv_pool_bytes=(len(knights) + 1) * max_ctx * kv_tok, flags=flags)
        adapters = {k.adapter: B200LocalLlmAdapter(f"b200://{device}", w["model"], k.name, max_new_tokens=nt, ignore_eos=True,
                                                   engine=eng) for k in knights}
        per_turn = []

        def on_turn(turn, adapter):
            g = adapter.last_generation
            per_turn.append((turn.round, g.n_prompt_tokens, g.prefix_cached_tokens, g.prefill_ms, g.decode_ms, g.decode_steps, turn.wall_ms))
        t0 = time.perf_counter()
        res = R.run_discussion(adapters, knights, "Should the cache layer move in front of the API gateway?", max_rounds=rounds,
                               rng=random.Random(1234), on_turn=on_turn)
        wall = (time.perf_counter() - t0) * 1e3
        eng.close()
        errs = [f"{t.knight} r{t.round}: {t.error}" for t in res.turns if t.error]
        if errs or len(per_turn) != rounds * len(knights):
            raise RuntimeError(f"discussion incomplete ({len(per_turn)} turns): {errs[:2]}")
        out[label] = {
            "wall_ms": wall, "turns": len(per_turn),
            "round_ms": [sum(t[6] for t in per_turn if t[0] == r) for r in range(1, rounds + 1)],
            "prompt_tokens": sum(t[1] for t in per_turn), "prefix_cached_tokens": sum(t[2] for t in per_turn),
            "hit_rate": sum(t[2] for t in per_turn) / max(1, sum(t[1] for t in per_turn)),
            "prefill_ms": sum(t[3] for t in per_turn), "decode_ms": sum(t[4] for t in per_turn),
            "decode_tok_per_s": sum(t[5] for t in per_turn) / (sum(t[4] for t in per_turn) * 1e-3),
            "last_prompt_tokens": per_turn[-1][1],
        }
    out["note"] = (f"{len(knights)} knights x {rounds} rounds, {nt} new tokens per turn, byte-level tokenizer, synthetic template of the "
                   "reference's size and placeholder layout; SEQ = what `roundtable discuss` issues.  Random weights emit ids outside the byte "
                   "range, each detokenised as '<t12345>' and read back by the byte-level tokenizer as ~9 tokens, so the transcript grows "
                   "~4.7k tokens per turn (about 8x a BPE server's): prompt and prefill figures are pessimistic by that factor")
    return out


def run_b200(args, w):
    from theroundtaible_b200 import engine as E
    D = Dist()
    rank, world = D.rank, D.world
    tp = 1 if args.replicas else (args.tp if args.tp > 0 else world)
    if tp > 1:
        assert world == tp, "tensor parallel needs one rank per GPU (torchrun --nproc-per-node N)"
    model = w["model"] + (f"@L{args.layers}" if args.layers else "")
    sampler = ClockSampler(D.local) if rank == 0 else None
    h = measure(D, E, w, model, tp, args.steps, args.warmup, e2e=True, seq=True, sampler=sampler)
    cfg = h["cfg"]

    extra = {"seq": h["seq"], "batch_round_ms": h["ms_per_step"]}
    quick = dict(steps=min(args.steps, 3), warmup=min(args.warmup, 2))
    if not args.no_extras and args.workload == "config3" and not args.layers:
        try:
            if world == 1:
                w2 = dict(WORKLOADS["config2"])
                m2 = measure(D, E, w2, w2["model"], 1, min(args.steps, 5), min(args.warmup, 3), e2e=True, seq=True)
                extra["config2"] = brief(m2, w2, w2["model"], 1)
            elif tp in (2, 4):
                w4 = dict(WORKLOADS["config4"])
                extra["config4"] = brief(measure(D, E, w4, w4["model"], tp, e2e=False, **quick), w4, w4["model"], tp)
            elif tp == 8:
                w5 = dict(WORKLOADS["config5"])
                extra["config5"] = brief(measure(D, E, w5, w5["model"], tp, e2e=False, **quick), w5, w5["model"], tp)
        except Exception as ex:  # noqa: BLE001
            extra["extras_failed"] = repr(ex)
    if not args.no_tp_check:
        try:
            extra["tp_check"] = tp_check(D, E, w, tp)
        except Exception as ex:  # noqa: BLE001
            extra["tp_check"] = {"ok": False, "failed": repr(ex)}

    if rank == 0:
        hbm = h["peak"]
        # the dominant kernel alone: the gate/up projection (fused SwiGLU epilogue) of one layer of THIS rank, re-streamed from HBM
        gu_ms = E.k_gemv_bench(2 * cfg.ffn // tp, cfg.hidden, w["knights"], iters=200, device=D.local, epi=2)
        gu_gbs = 2 * (cfg.ffn // tp) * cfg.hidden * 2 / (gu_ms * 1e-3) / 1e9
        nw = 16 if cfg.hidden >= 8192 else 8
        xm = 2 if 5 <= w["knights"] <= 8 and os.environ.get("RT_DEC_F16", "1") != "0" else (1 if w["knights"] <= 4 else 0)
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            try:
                from oracle.build import build as obuild
                try:
                    obuild()
                except Exception:  # noqa: BLE001
                    pass
                _, r = cpu_round(w, args.cpu_layers, args.cpu_new_tokens)
                cpu = {"value": r["decode_tok_per_s"], "unit": "tok/s", "cores": r["cores"], "kind": "port", "sample": r["sample"],
                       "extrapolated": r["extrapolated"], "round_ms": r["round_ms"], "e2e_tok_per_s": r["e2e_tok_per_s"]}
            except Exception as ex:  # noqa: BLE001
                cpu = {"value": None, "unit": "tok/s", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {ex!r}"}
        if not args.no_discuss and world == 1 and not args.no_extras:
            try:
                w2 = dict(WORKLOADS["config2"])
                extra["discuss"] = discuss_leg(w2, w2["model"], D.local)
            except Exception as ex:  # noqa: BLE001
                extra["discuss"] = {"failed": repr(ex)}
        # DRAM traffic of one decode step as ncu saw it: a committed capture of the same workload, NOT measured in this run
        traffic, traffic_src, dom_traffic, dom_traffic_src = None, None, None, None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            tr = tj.get(f"{args.workload}_tp{tp}")
            if tr and not args.layers:
                traffic, traffic_src = tr["dram_bytes_per_step"], tr["source"]
            dk = tj.get(f"dominant_kernel_{args.workload}")
            if dk and tp == 1 and not args.layers:
                dom_traffic, dom_traffic_src = dk["dram_bytes_per_launch"], dk["source"]
        except Exception:  # noqa: BLE001
            pass
        par = f"tp{tp}" if tp > 1 else (f"replicas x{world}" if world > 1 else "single GPU")
        conf = workload_config(args.workload, w, model, par)
        conf.update({"prefill_ms_per_round": h["prefill_ms"], "decode_ms_per_token_step": h["step_ms"],
                     "numerics": "bf16 weights (the prefill GEMMs and decode steps of 5-8 slots stream their exact fp16 image) / bf16 Q / KV, "
                                 "fp16 projection inputs, fp32 accumulate + residual + softmax"})
        line = {
            "metric": "decode_tok_per_s", "value": h["value"], "unit": "tok/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": h["ms_per_step"], "higher_is_better": True,
            "scaling": "strong" if tp > 1 or world == 1 else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": conf,
            "roofline": {"bound": "hbm", "achieved": h["achieved_gbs"], "peak": hbm, "unit": "GB/s", "frac": h["frac"],
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": h["peak_src"],
                         "of": "decode step on one GPU (one CUDA-graph launch), algorithmic bytes / CUDA-event time",
                         "algorithmic_bytes_per_launch": h["bytes_step"],
                         "dominant_kernel": {"name": f"gemv_kernel<1,MT,EPI_SWIGLU,{nw},{xm}> (gate/up projection of one layer; "
                                                     f"{'fp16 weight twins x fp16 inputs' if xm == 2 else 'bf16 weights x bf16 hi + lo of the fp16 inputs'})",
                                             "timed": "alone, CUDA events, weights rotated through > 3x L2", "ms": gu_ms,
                                             "algorithmic_bytes_per_launch": 2 * (cfg.ffn // tp) * cfg.hidden * 2,
                                             "achieved": gu_gbs, "frac": gu_gbs / hbm, "traffic": dom_traffic,
                                             "traffic_source": dom_traffic_src}},
            "cpu_baseline": cpu, "e2e": h["e2e"], "gpu_launches": h["launches"], "clocks": h.get("clocks"),
            "prefill": h["prefill"], "extra": extra,
        }
        print(json.dumps(line), flush=True)
    D.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-discuss", action="store_true", help="skip the 5-round SEQ discussion leg (extra.discuss)")
    ap.add_argument("--no-extras", action="store_true", help="headline workload only")
    ap.add_argument("--no-tp-check", action="store_true")
    ap.add_argument("--tp", type=int, default=0, help="tensor-parallel degree; default = number of ranks")
    ap.add_argument("--replicas", action="store_true", help="N independent replicas instead of tensor parallel over the N ranks")
    ap.add_argument("--layers", type=int, default=0, help="truncate the model depth (stated in config.model)")
    ap.add_argument("--cpu-layers", type=int, default=0,
                    help="CPU arm: decoder layers run at true dimensions, the rest extrapolated (0 = all for models under 3B parameters, else 1)")
    ap.add_argument("--cpu-new-tokens", type=int, default=8, help="CPU arm: decode steps run per round (rest extrapolated)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    w = dict(WORKLOADS[args.workload])
    if args.impl == "reference":
        run_reference(args, w)
    else:
        run_b200(args, w)


if __name__ == "__main__":
    main()
