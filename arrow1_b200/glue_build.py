"""Builds arrow_glue/ (the Arrow FunctionRegistry glue + its GPU parity driver) where the reference
headers exist; on the GPU box the prebuilt files shipped with the snapshot are used."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GLUE = os.path.join(ROOT, "arrow_glue")
TEST_BIN = os.path.join(GLUE, "test_glue")
ARCHIVE = os.path.join(GLUE, "libarrow1_b200_arrow.a")


def build(verbose: bool = False) -> str:
    if not os.path.isdir("/root/reference/cpp/src/arrow"):
        return TEST_BIN
    srcs = [os.path.join(GLUE, f) for f in ("b200_arrow.cc", "b200_arrow.h", "test_glue.cc", "build.sh")]
    srcs.append(os.path.join(ROOT, "include", "arrow1_b200.h"))
    if all(os.path.exists(p) for p in (TEST_BIN, ARCHIVE)):
        t = min(os.path.getmtime(TEST_BIN), os.path.getmtime(ARCHIVE))
        if all(os.path.getmtime(s) <= t for s in srcs):
            return TEST_BIN
    r = subprocess.run(["bash", os.path.join(GLUE, "build.sh")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("arrow_glue build failed:\n" + r.stdout[-6000:])
    if verbose:
        print(r.stdout)
    return TEST_BIN


if __name__ == "__main__":
    print(build(verbose=True))
