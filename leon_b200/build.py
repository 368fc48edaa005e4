"""Builds libleon_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC_DIR = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libleon_b200.so")
CLI = os.path.join(HERE, "bin", "leon_b200")
CLI_SRC = os.path.join(HERE, "cli", "main.cpp")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def sources():
    return sorted(os.path.join(SRC_DIR, f) for f in os.listdir(SRC_DIR))


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    hdr = os.path.join(os.path.dirname(HERE), "include", "leon_b200.h")
    if not os.path.exists(CLI) or os.path.getmtime(CLI_SRC) > os.path.getmtime(CLI):
        return True
    return any(os.path.getmtime(p) > t for p in sources() + [hdr])


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
        ["-I", SRC_DIR, "-o", LIB, os.path.join(SRC_DIR, "leon_b200.cu"), os.path.join(SRC_DIR, "container.cpp"),
         os.path.join(SRC_DIR, "headers.cpp"), os.path.join(SRC_DIR, "fileio.cpp"),
         os.path.join(SRC_DIR, "container_read.cpp"), os.path.join(SRC_DIR, "decode.cu"),
         os.path.join(SRC_DIR, "decompress.cpp"), "-lz"]
    subprocess.run(cmd, check=True)
    # the command line tool (host only), finds the library next to it
    os.makedirs(os.path.dirname(CLI), exist_ok=True)
    subprocess.run([os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-o", CLI, CLI_SRC, "-L", HERE, "-lleon_b200",
                    "-Wl,-rpath,$ORIGIN/.."], check=True)
    # the same program under the reference's name: `leon -file X -c|-d`
    alias = os.path.join(os.path.dirname(CLI), "leon")
    if os.path.lexists(alias):
        os.remove(alias)
    os.link(CLI, alias)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
