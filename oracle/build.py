"""Build recipe for the oracle's C/C++ restatements (gcc only; outputs to oracle/_build/).
TEST INFRASTRUCTURE.  Run: python -m oracle.build"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_build")
TARGETS = {
    "librans64_ref.so": (["gcc", "-O2", "-shared", "-fPIC", "-std=c11"], ["csrc/rans64_ref.c"]),
    "libsparse_conv_ref.so": (["g++", "-O3", "-march=native", "-fopenmp", "-shared", "-fPIC", "-std=c++17"],
                              ["csrc/sparse_conv_ref.cpp"]),
}


def build(force=False):
    os.makedirs(OUT, exist_ok=True)
    for name, (cmd, srcs) in TARGETS.items():
        dst = os.path.join(OUT, name)
        src_paths = [os.path.join(HERE, s) for s in srcs]
        if not force and os.path.exists(dst) and all(os.path.getmtime(dst) >= os.path.getmtime(s) for s in src_paths):
            continue
        subprocess.check_call(cmd + src_paths + ["-o", dst])
    return OUT


def lib_path(name):
    p = os.path.join(OUT, name)
    if not os.path.exists(p):
        build()
    return p


if __name__ == "__main__":
    build(force=True)
    print("oracle built ->", OUT)
