"""Build recipe for the CPU oracle (test infrastructure; see flat_l2_oracle.c header).

`python oracle/build.py` -> oracle/liboracle.so (git-ignored, travels to the GPU box).
There is no `oracle/_ref`: the reference's arithmetic is the third-party faiss-cpu wheel,
whose source is not under /root/reference and which cannot be installed here, so there is
nothing of the reference's to compile ("Oracle").
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "flat_l2_oracle.c")
OUT = os.path.join(HERE, "liboracle.so")


def build(force: bool = False) -> str:
    if (not force and os.path.exists(OUT)
            and os.path.getmtime(OUT) >= os.path.getmtime(SRC)):
        return OUT
    cmd = ["gcc", "-O3", "-fopenmp", "-shared", "-fPIC", "-fvisibility=hidden",
           "-fno-math-errno", "-o", OUT, SRC, "-lm"]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
