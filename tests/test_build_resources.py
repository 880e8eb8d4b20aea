"""Resource usage of the built kernels (cuobjdump on the in-tree library; no GPU needed).

The decode projections are sized by their registers: an 8-warp CTA must stay at <= 128 registers per thread so that TWO fit on
an SM (twice the weight loads in flight), the three-per-SM form at <= 85, and none may spill.  A launch-bounds slip that let
the compiler take 134-150 registers cost every configuration 5-22 % during round 2 (profiles/r2_summary.md 7) -- hence a test."""
import os
import re
import shutil
import subprocess

import pytest

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "theroundtaible_b200", "_lib", "librt_engine.so")


def _resource_usage():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not available")
    if not os.path.exists(LIB):
        from theroundtaible_b200.build import build
        build()
    out = subprocess.run([exe, "--dump-resource-usage", LIB], capture_output=True, text=True, check=True).stdout
    res = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+)", out):
        res[m.group(1)] = (int(m.group(2)), int(m.group(3)), int(m.group(4)))
    assert res, "no kernels found in the library"
    return res


def test_projection_kernels_keep_their_occupancy():
    res = _resource_usage()
    gemv = {k: v for k, v in res.items() if k.startswith("_Z11gemv_kernelI")}
    assert len(gemv) > 100
    for name, (reg, stack, _) in gemv.items():
        # template arguments <NT, MT, EPI, NW, XM, CL, OC> as mangled: ...ILi1ELi2ELi2ELi8ELi2ELi1ELi1EE...
        args = [int(x) for x in re.findall(r"Li(\d+)E", name)]
        assert len(args) == 7, name
        nw, oc = args[3], args[6]
        assert stack == 0, f"{name} spills ({stack} bytes)"
        if oc == 3:
            assert reg * 32 * nw * 3 <= 65536, f"{name}: {reg} registers, three CTAs per SM no longer fit"
        elif nw == 8:
            assert reg <= 128, f"{name}: {reg} registers, two 8-warp CTAs per SM no longer fit"
        else:
            assert reg <= 128, f"{name}: {reg} registers, a 16-warp CTA no longer fits"


def test_tensor_memory_attention_kernels_fit_one_cta_per_sm():
    res = _resource_usage()
    names = [k for k in res if "attn_decode_tc5_kernel" in k or "attn_prefill_tc5b_kernel" in k or "gemm_tc5_kernel" in k]
    assert names
    for name in names:
        reg, _, _ = res[name]
        assert reg <= 255 and reg > 0, name
    for name in (k for k in res if "attn_decode_tc5_kernel" in k):
        assert res[name][0] * 320 <= 65536, f"{name}: {res[name][0]} registers x 320 threads exceed the register file"
