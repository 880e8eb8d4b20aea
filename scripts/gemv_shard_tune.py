"""Projection micro-bench (kernel alone, re-streamed from HBM, CUDA events) over the per-layer shapes of tensor-parallel
SHARDS, under the dispatch switches of engine.cu (rows per CTA, warps, persistent walking of small row groups).  One process
per setting (the switches are read once).   usage: gemv_shard_tune.py            (driver)
                                                    gemv_shard_tune.py one <T> <name:N:K:epi> ...   (worker)"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

SHARDS = {   # (T, [(name, N, K, epi)])
    "70b/tp8": (8, [("qkv", 1280, 8192, 0), ("o", 8192, 1024, 0), ("gu", 7168, 8192, 2), ("down", 8192, 3584, 0)]),
    "32b/tp4": (8, [("qkv", 1792, 5120, 0), ("o", 5120, 1280, 0), ("gu", 13824, 5120, 2), ("down", 5120, 6912, 0)]),
    "8b/tp1": (5, [("qkv", 6144, 4096, 0), ("o", 4096, 4096, 0), ("gu", 28672, 4096, 2), ("down", 4096, 14336, 0)]),
    "8b/tp2": (5, [("qkv", 3072, 4096, 0), ("o", 4096, 2048, 0), ("gu", 14336, 4096, 2), ("down", 4096, 7168, 0)]),
    "8b/tp4": (5, [("qkv", 1536, 4096, 0), ("o", 4096, 1024, 0), ("gu", 7168, 4096, 2), ("down", 4096, 3584, 0)]),
    "8b/tp8": (5, [("qkv", 768, 4096, 0), ("o", 4096, 512, 0), ("gu", 3584, 4096, 2), ("down", 4096, 1792, 0)]),
}
SETTINGS = [{}, {"RT_GEMV_MT": "1"}, {"RT_GEMV_MT": "2"}, {"RT_GEMV_MT": "4"}, {"RT_GEMV_NW": "8"}, {"RT_GEMV_NW": "16"},
            {"RT_GEMV_PMODE": "1"}, {"RT_GEMV_PMODE": "2"}, {"RT_GEMV_NW": "16", "RT_GEMV_CAP16": "2"}]

if len(sys.argv) > 1 and sys.argv[1] == "one":
    from theroundtaible_b200 import engine as E
    T = int(sys.argv[2])
    out = []
    for spec in sys.argv[3:]:
        name, N, K, epi = spec.split(":")
        N, K, epi = int(N), int(K), int(epi)
        ms = E.k_gemv_bench(N, K, T, iters=300, epi=epi)
        out.append(f"{name} {ms * 1e3:5.1f}us {N * K * 2 / ms / 1e6:5.0f}GB/s")
    print(" | ".join(out), flush=True)
else:
    only = sys.argv[1:] or list(SHARDS)
    for shard in only:
        T, shapes = SHARDS[shard]
        for st in SETTINGS:
            env = dict(os.environ, **st)
            r = subprocess.run([sys.executable, __file__, "one", str(T)] + [f"{n}:{N}:{K}:{e}" for n, N, K, e in shapes],
                               env=env, capture_output=True, text=True)
            tag = " ".join(f"{k[8:]}={v}" for k, v in st.items()) or "default"
            print(f"{shard:8s} {tag:18s} {r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-200:]}", flush=True)
