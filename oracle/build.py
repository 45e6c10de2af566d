"""Build recipe for the CPU oracle (test infrastructure, not product).

  python oracle/build.py        ->  oracle/libkmeans_oracle.so

The reference (schusto/k-means-demo) has no compiled sources and no Lloyd path, so there
is nothing to compile into oracle/_ref/ — "unbuildable: nothing to build" (DESIGN.md §3).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "kmeans_oracle.c")
OUT = os.path.join(HERE, "libkmeans_oracle.so")

# -ffp-contract=off: no FMA contraction, so doubles round exactly as a JS engine's would.
CFLAGS = ["-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-shared", "-fPIC", "-std=c11"]


def build(force: bool = False) -> str:
    if (not force and os.path.exists(OUT)
            and os.path.getmtime(OUT) >= os.path.getmtime(SRC)):
        return OUT
    cmd = ["gcc", *CFLAGS, "-o", OUT, SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
