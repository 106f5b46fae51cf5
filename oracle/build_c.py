"""ORACLE (test infrastructure only): builds the plain-C pieces of the oracle (oracle/slic_c.c) with gcc into
oracle/_build/liboracle_c.so.  Called by __graft_entry__.build() and, lazily, by oracle.slic when the library is
missing (gcc is part of the image, here and on the GPU box)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_build", "liboracle_c.so")
SRCS = [os.path.join(HERE, "slic_c.c")]


def build(force=False):
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(s) for s in SRCS):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    # -ffp-contract=off: every float64 operation rounds on its own, as in the cython / numpy loops restated
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", OUT] + SRCS + ["-lm"])
    return OUT


if __name__ == "__main__":
    print(build(force=True))
