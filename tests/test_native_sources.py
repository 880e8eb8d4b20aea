"""The Node-API addon and the TypeScript adapter class cannot be built here -- no Node, no
node_api.h, no tsc -- but they are files a compiler has seen: the C source must pass `gcc -fsyntax-only -Wall -Werror`
against a declarations-only stand-in of the N-API functions it calls and the REAL include/rt_engine.h, so every use of
the C-ABI in it (struct fields, argument order) is type-checked; the TypeScript class must keep the reference adapter's
surface."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
def test_addon_type_checks_against_the_abi_header():
    r = subprocess.run(["gcc", "-std=c99", "-fsyntax-only", "-Wall", "-Wextra", "-Werror", "-DNODE_GYP_MODULE_NAME=rt_addon",
                        "-I", os.path.join(ROOT, "tests", "stubs"), "-I", os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "native", "rt_addon.c")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_addon_calls_only_functions_the_library_exports():
    src = open(os.path.join(ROOT, "native", "rt_addon.c")).read()
    hdr = open(os.path.join(ROOT, "include", "rt_engine.h")).read()
    declared = set(re.findall(r"RT_API\s+[\w\s\*]+?\b(rt_\w+)\s*\(", hdr))
    used = set(re.findall(r"\b(rt_[a-z_0-9]+)\s*\(", src))
    assert used and used <= declared, used - declared


def test_typescript_adapter_keeps_the_reference_surface():
    ts = open(os.path.join(ROOT, "native", "local-llm-b200.ts")).read()
    assert "extends BaseAdapter" in ts and "readonly name: string" in ts                       # base.ts:10-12
    for sig in ("async isAvailable(): Promise<boolean>", "override getMaxSourceChars(): number | undefined",
                "async execute(prompt: string, timeoutMs?: number): Promise<string>"):          # base.ts:16-24
        assert sig in ts, sig
    assert "constructor(endpoint: string, model: string, name: string, _source?: string, timeoutMs: number = 120_000)" in ts
    assert "classifyError(error, this.name)" in ts and "Local LLM returned empty response" in ts
    assert ts.count("{") == ts.count("}") and ts.count("(") == ts.count(")")
    for fn in ("create", "generate", "contextLength", "destroy"):                              # what rt_addon.c exports
        assert f'"{fn}"' in open(os.path.join(ROOT, "native", "rt_addon.c")).read() and f"native.{fn}(" in ts or fn == "destroy"
