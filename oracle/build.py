"""Build recipe for the C restatement (oracle/explainn_oracle.c -> oracle/_build/liborc.so).

TEST INFRASTRUCTURE ONLY.  The reference itself is pure Python (no C/C++ sources), so there is
no `oracle/_ref` binary to build; the reference is exercised by tests/golden/make_golden.py in
the build container instead and its outputs are committed as fixtures.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "explainn_oracle.c")
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "liborc.so")


def build(force=False):
    os.makedirs(OUT_DIR, exist_ok=True)
    if (not force and os.path.exists(LIB)
            and os.path.getmtime(LIB) >= os.path.getmtime(SRC)):
        return LIB
    cmd = ["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared",
           "-o", LIB, SRC, "-lm"]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
