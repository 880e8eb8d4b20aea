#!/usr/bin/env python
"""Benchmark of the local-model generation path (BASELINE.json metric: decode tok/s + full-round latency).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload config2|config3]

A STEP is one round of the roundtable in BATCH mode: the N knight prompts of the round (shared prefix +
per-knight suffix, synthetic token ids) are prefilled and every knight generates `new_tokens` tokens
greedily.  `value` is whole-job DECODE tok/s from the engine's CUDA events on its own stream (state
already resident in HBM); `ms_per_step` is the full-round latency; `e2e` is the same round driven
through the adapter API with HOST strings (tokenise -> H2D -> generate -> D2H -> detokenise).  `extra.seq`
reports the SEQ mode the reference orchestrator really runs (one knight per call).

With --gpus N > 1 (torchrun) every rank runs an independent replica of the round on its own GPU (weak
scaling; 1B/8B models do not shard -- DESIGN.md "multi-GPU").

--impl reference times the CPU oracle server (the reference delegates to an external server that cannot
be installed here; DESIGN.md "oracle") on all host cores, on a bounded sample of the same workload.
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

WORKLOADS = {
    # BASELINE.json configs[1]: 3 knights, Llama-3.2-1B bf16, shared system-prompt prefix, greedy
    "config2": dict(model="llama-3.2-1b", knights=3, shared=1024, own=64, new_tokens=512, max_ctx=2048, seed_idx=2),
    # BASELINE.json configs[2]: 5 knights, Llama-3.1-8B, 8k shared transcript prefix
    "config3": dict(model="llama-3.1-8b", knights=5, shared=8192, own=128, new_tokens=512, max_ctx=9216, seed_idx=3),
    # BASELINE.json configs[3]/[4]: tensor parallel (run with --tp N under torchrun); layer count may be truncated with --layers
    "config4": dict(model="qwen2.5-32b", knights=8, shared=2048, own=128, new_tokens=256, max_ctx=2560, seed_idx=4),
    "config5": dict(model="llama-3.1-70b", knights=8, shared=16384, own=128, new_tokens=256, max_ctx=16896, seed_idx=5),
    # smoke-sized
    "tiny": dict(model="tiny-llama", knights=3, shared=64, own=8, new_tokens=32, max_ctx=256, seed_idx=0),
}
P_DEC = {"llama-3.2-1b": None, "llama-3.1-8b": None}


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
        # median over samples taken under load (upper half: idle samples between steps would drag it down)
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# =====================================================================================================
def run_reference(args, w):
    """CPU arm: the oracle server on all host cores, bounded sample of the same round."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    import torch
    from oracle.build import build as obuild
    try:
        obuild()
    except Exception:  # noqa: BLE001
        pass
    from oracle.server import OracleServer
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sample = dict(shared=min(w["shared"], 96), own=min(w["own"], 16), new_tokens=min(w["new_tokens"], 12))
    srv = OracleServer(w["model"], seed=0, max_pos=sample["shared"] + sample["own"] + sample["new_tokens"] + 8)
    ws = dict(w, **sample)
    prompts = synth_prompts(ws, srv.m.a.vocab)
    shared_pages = (sample["shared"] // 16) * 16
    dec_s, tot_s, toks = 0.0, 0.0, 0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        _, t = srv.complete_batch_ids(prompts, sample["new_tokens"], shared=shared_pages)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            dec_s += t["decode_s"]
            tot_s += dt
            toks += w["knights"] * t["decode_steps"]
    value = toks / dec_s
    desc = (f"{w['knights']} knights, {sample['shared']} shared + {sample['own']} own prompt tokens, "
            f"{sample['new_tokens']} new tokens per knight, fp32 compute over bf16 weights, batched decode")
    line = {"impl": "reference", "metric": "decode_tok_per_s", "value": value, "unit": "tok/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {w['knights']} knights, {w['model']}, BATCH round (bounded CPU sample)",
                       "model_weights": "seeded random-init, bf16 values", "sample": desc},
            "cpu_baseline": {"value": value, "unit": "tok/s", "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": value, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_baseline_leg(w):
    """Bounded CPU sample of the same workload on rank 0 (reported beside the GPU number, not a target)."""
    import torch
    from oracle.build import build as obuild
    try:
        obuild()
    except Exception:  # noqa: BLE001
        pass
    from oracle.server import OracleServer
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sample = dict(shared=min(w["shared"], 96), own=min(w["own"], 16), new_tokens=min(w["new_tokens"], 12))
    srv = OracleServer(w["model"], seed=0, max_pos=sample["shared"] + sample["own"] + sample["new_tokens"] + 8)
    prompts = synth_prompts(dict(w, **sample), srv.m.a.vocab)
    _, t = srv.complete_batch_ids(prompts, sample["new_tokens"], shared=(sample["shared"] // 16) * 16)
    return {"value": w["knights"] * t["decode_steps"] / t["decode_s"], "unit": "tok/s", "cores": cores, "kind": "port",
            "sample": (f"{w['knights']} knights, {sample['shared']}+{sample['own']} prompt tokens, {sample['new_tokens']} new "
                       f"tokens each on {w['model']}; oracle port, fp32 compute, torch threads={cores}; "
                       f"prefill {t['prefill_s']:.2f}s decode {t['decode_s']:.2f}s")}


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


# =====================================================================================================
def run_b200(args, w):
    import numpy as np
    import torch
    from theroundtaible_b200 import B200LocalLlmAdapter
    from theroundtaible_b200 import engine as E

    rank, world, local = dist_env()
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this backend has no CPU fallback)")

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    tp = args.tp if args.tp > 1 else 1
    model = w["model"] + (f"@L{args.layers}" if args.layers else "")
    if tp > 1:
        assert world == tp, "--tp N needs torchrun with N ranks"
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(E.nccl_unique_id()), dtype=torch.uint8))
        torch.distributed.broadcast(idt, 0)
        eng = E.Engine(model, seed=0, device=local, max_batch=max(w["knights"], 1), max_ctx=w["max_ctx"], tp_rank=rank,
                       tp_size=tp, nccl_id=bytes(idt.cpu().numpy().tobytes()))
    else:
        eng = E.Engine(model, seed=0, device=local, max_batch=max(w["knights"], 1), max_ctx=w["max_ctx"])
    replicas = 1 if tp > 1 else world
    prompts = synth_prompts(w, eng.vocab, 0 if tp > 1 else rank)
    texts = synth_text_prompts(w, 0 if tp > 1 else rank)
    adapter = B200LocalLlmAdapter(f"b200://{local}", w["model"], "bench", max_new_tokens=w["new_tokens"], ignore_eos=True,
                                  engine=eng)
    nt = w["new_tokens"]

    def step():
        return eng.generate_ids(prompts, max_new_tokens=nt, ignore_eos=True)

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    t0 = time.perf_counter()
    dec_ms = pre_ms = 0.0
    toks = launches = steps_dec = 0
    for _ in range(args.steps):
        gens = step()
        dec_ms += gens[0].decode_ms
        pre_ms += gens[0].prefill_ms
        steps_dec += gens[0].decode_steps
        toks += sum(g.decode_steps for g in gens)
        launches += gens[0].gpu_launches
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3

    # ---- e2e: the adapter API with host strings ----
    for _ in range(min(args.warmup, 2)):
        adapter.execute_round(texts)
    barrier()
    e0 = time.perf_counter()
    e2e_tokens = 0
    for _ in range(args.steps):
        outs = adapter.execute_round(texts)
        e2e_tokens += len(outs) * nt
    barrier()
    e2e_ms = (time.perf_counter() - e0) * 1e3
    clocks = sampler.stop() if rank == 0 else None

    # ---- SEQ mode (what orchestrator.ts:361-442 does): one knight per call, B = 1 ----
    seq_dec_ms, seq_tok = 0.0, 0
    s0 = time.perf_counter()
    for p in prompts:
        g = eng.generate_ids([p], max_new_tokens=nt, ignore_eos=True)[0]
        seq_dec_ms += g.decode_ms
        seq_tok += g.decode_steps
    seq_round_ms = (time.perf_counter() - s0) * 1e3

    # ---- max over ranks ----
    stats = torch.tensor([dec_ms, wall_ms, e2e_ms, pre_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        torch.distributed.all_reduce(stats, op=torch.distributed.ReduceOp.MAX)
    dec_ms_m, wall_ms_m, e2e_ms_m, pre_ms_m = [float(v) for v in stats.tolist()]

    if rank == 0:
        info = eng.info
        peak, peak_src = measured_peak_hbm()
        # algorithmic bytes of one decode step: weights once + shared prefix KV once + own KV per knight
        mean_own_ctx = w["own"] + nt / 2.0
        shared_pages = (w["shared"] // 16) * 16
        kv_b = info.kv_bytes_per_token * (shared_pages + w["knights"] * (w["shared"] - shared_pages + mean_own_ctx))
        bytes_step = info.weight_bytes_per_step + kv_b
        step_ms = dec_ms_m / max(steps_dec, 1)
        achieved = bytes_step / (step_ms * 1e-3) / 1e9
        # the dominant kernel alone: the gate/up projection stream of one layer, re-streamed from HBM
        cfg = eng.cfg
        # prefill: tensor-core path.  FLOPs = 2 * projection weights * tokens + causal attention (QK^T and PV)
        pf_tokens = shared_pages + w["knights"] * (w["shared"] - shared_pages + w["own"])
        per_layer_w = (cfg.hidden * (cfg.n_q + 2 * cfg.n_kv) * cfg.head_dim + cfg.n_q * cfg.head_dim * cfg.hidden + 3 * cfg.hidden * cfg.ffn) // tp
        attn_pairs = shared_pages * (shared_pages + 1) / 2 + w["knights"] * sum(range(shared_pages + 1, w["shared"] + w["own"] + 1))
        pf_flops = cfg.layers * (2.0 * per_layer_w * pf_tokens + 4.0 * (cfg.n_q // tp) * cfg.head_dim * attn_pairs)
        pf_ms = pre_ms_m / args.steps
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            tf_burst, tf_sust = float(peaks["bf16_tflops"]), float(peaks["bf16_tflops_sustained"])
        except Exception:  # noqa: BLE001
            tf_burst, tf_sust = 1590.0, 1400.0
        pf_tflops = pf_flops / (pf_ms * 1e-3) / 1e12
        gu_ms = E.k_gemv_bench(2 * cfg.ffn // tp, cfg.hidden, w["knights"], iters=200, device=local)
        gu_gbs = 2 * (cfg.ffn // tp) * cfg.hidden * 2 / (gu_ms * 1e-3) / 1e9
        cpu = None
        if not args.no_cpu_baseline and tp == 1 and info.n_params < 3e9:
            try:
                cpu = cpu_baseline_leg(w)
            except Exception as ex:  # noqa: BLE001
                cpu = {"value": None, "unit": "tok/s", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {ex}"}
        discuss = None
        if not args.no_discuss and tp == 1 and world == 1 and info.n_params < 3e9:
            try:
                discuss = discuss_leg(w, model, local)
            except Exception as ex:  # noqa: BLE001
                discuss = {"failed": repr(ex)}
        # DRAM traffic of one step as ncu saw it (committed capture of the same workload), next to the algorithmic bytes
        traffic = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r1c_traffic.json"))).get(args.workload)
            if tr and tp == 1 and not args.layers:
                traffic = tr["dram_read_bytes_per_step"]
        except Exception:  # noqa: BLE001
            pass
        h2d = sum(len(t.encode()) + 3 for t in texts) * 4 + 4 * 96 * ((sum(len(t.encode()) + 3 for t in texts) + 31) // 32)
        d2h = w["knights"] * nt * 4
        line = {
            "metric": "decode_tok_per_s", "value": replicas * toks / (dec_ms_m * 1e-3), "unit": "tok/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": wall_ms_m / args.steps, "higher_is_better": True,
            "scaling": "strong" if tp > 1 else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": (f"{args.workload}: {w['knights']} knights, {w['model']} bf16, {w['shared']} shared + "
                                    f"{w['own']} own prompt tokens, {nt} new tokens per knight, greedy, BATCH round"),
                       "weights": "seeded random-init (Irwin-Hall approx of N(0,0.02)), bf16", "l2": "weights (2.47 GB+) exceed the 126 MB L2",
                       "parallelism": (f"tp{tp}" if tp > 1 else (f"replicas x{world}" if world > 1 else "single GPU")),
                       "model": model,
                       "prefill_ms_per_round": pre_ms_m / args.steps, "decode_ms_per_token_step": step_ms},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "of": "decode step (one CUDA-graph launch)",
                         "algorithmic_bytes_per_launch": bytes_step,
                         "dominant_kernel": {"name": "gemv_tc_kernel (gate/up projection)", "ms": gu_ms, "achieved": gu_gbs,
                                             "frac": gu_gbs / peak}},
            "cpu_baseline": cpu,
            "e2e": {"value": replicas * e2e_tokens / (e2e_ms_m * 1e-3), "unit": "tok/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_round": e2e_ms_m / args.steps,
                    "note": "generated tokens / wall time of adapter.execute_round (tokenise, prefill, decode, detokenise)"},
            "gpu_launches": launches, "clocks": clocks,
            "prefill": {"tokens_prefilled": pf_tokens, "ms": pf_ms, "tok_per_s": pf_tokens / (pf_ms * 1e-3), "tflops": pf_tflops,
                        "bound": "tensor", "frac_of_measured_burst": pf_tflops / tf_burst, "frac_of_measured_sustained": pf_tflops / tf_sust,
                        "note": "whole prefill phase (tcgen05 GEMMs + flash attention + norms), algorithmic FLOPs / CUDA-event time"},
            "extra": {"seq": {"decode_tok_per_s": seq_tok / (seq_dec_ms * 1e-3), "round_ms": seq_round_ms,
                              "note": "SEQ = one knight per call (B=1), what orchestrator.ts:361-442 issues"},
                      "batch_round_ms": wall_ms_m / args.steps, "discuss": discuss},
        }
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        torch.distributed.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-discuss", action="store_true", help="skip the 5-round SEQ discussion leg (extra.discuss)")
    ap.add_argument("--tp", type=int, default=1, help="tensor-parallel degree (= torchrun ranks); default: replicas")
    ap.add_argument("--layers", type=int, default=0, help="truncate the model depth (stated in config.model)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    w = dict(WORKLOADS[args.workload])
    if args.impl == "reference":
        run_reference(args, w)
    else:
        run_b200(args, w)


if __name__ == "__main__":
    main()
