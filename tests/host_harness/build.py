"""TEST INFRASTRUCTURE ONLY -- builds tests/host_harness/libaslr_hostmath.so with g++."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
LIB = os.path.join(HERE, "libaslr_hostmath.so")


def build(force=False):
    srcs = [os.path.join(HERE, "harness.cpp")]
    deps = srcs + [os.path.join(ROOT, "aslr_to_b200", "csrc", f) for f in os.listdir(os.path.join(ROOT, "aslr_to_b200", "csrc"))]
    deps.append(os.path.join(ROOT, "include", "aslr_b200.h"))
    if not force and os.path.isfile(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps):
        return LIB
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", "-I", os.path.join(ROOT, "include"), "-I",
           os.path.join(ROOT, "aslr_to_b200", "csrc"), "-o", LIB] + srcs
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
