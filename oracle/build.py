"""Build the oracle's C helpers into oracle/_build/ (TEST INFRASTRUCTURE; see oracle/__init__.py).

The reference (/root/reference) is TypeScript with no numeric code and its arithmetic lives in an
un-vendored external server, so there is nothing to compile into oracle/_ref/: the reference arm of the
benchmark is the oracle port itself (DESIGN.md "oracle")."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_build")
LIB = os.path.join(OUT, "liboracle_wgen.so")


def build(force: bool = False) -> str:
    os.makedirs(OUT, exist_ok=True)
    src = os.path.join(HERE, "csrc", "wgen.c")
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(src):
        return LIB
    subprocess.run(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared", "-o", LIB, src], check=True)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
