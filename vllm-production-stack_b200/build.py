"""Build libb200router.so (nvcc, sm_100a) and the host-side packer libb200pack.so (g++)
in-tree -- no JIT cache, no torch."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "b200_router", "libb200router.so")
PACK_OUT = os.path.join(HERE, "b200_router", "libb200pack.so")
SOURCES = ["prefix.cu", "multigpu.cu", "session.cu"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared", "-cudart", "shared",
         "-I", os.path.join(ROOT, "include"), "-I", CSRC]


def nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def needs_build():
    if not os.path.exists(OUT) or not os.path.exists(PACK_OUT):
        return True
    t = min(os.path.getmtime(OUT), os.path.getmtime(PACK_OUT))
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + \
           [os.path.join(ROOT, "include", "b200_router.h"), os.path.join(ROOT, "include", "b200_pack.h"),
            os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    cmd = [nvcc()] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          [os.path.join(CSRC, s) for s in SOURCES] + ["-o", OUT]
    subprocess.check_call(cmd)
    build_packer()
    return OUT


def build_packer():
    """The CPython-facing packer: plain C++, needs only Python.h."""
    import sysconfig
    cmd = [os.environ.get("CXX", "g++"), "-O3", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-pthread",
           "-I", sysconfig.get_paths()["include"], "-I", os.path.join(ROOT, "include"),
           os.path.join(CSRC, "packer.cpp"), "-o", PACK_OUT]
    subprocess.check_call(cmd)
    return PACK_OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
