"""CPU check of the device deflate stage (leon_b200/csrc/deflate.cuh): tools/deflate_sim.cpp compiles the SAME per-thread
phase functions the kernels call (k_deflate_count / _codes / _emit) for the host and runs them thread after thread;
every stream must inflate with zlib to the input, and stay close to zlib's own run-length strategy in size.  The
kernels themselves are covered on the GPU by tests/test_gpu_container.py."""
import os
import subprocess
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def sim(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("dfsim") / "deflate_sim")
    subprocess.run(["g++", "-O2", "-std=c++17", "-w", "-o", exe, os.path.join(ROOT, "tools", "deflate_sim.cpp")], check=True)
    return exe


def _texts():
    rng = np.random.default_rng(1)
    fib = [1, 1, 2, 3, 5, 8, 13, 21, 34, 55, 89, 144, 233, 377, 610, 987, 1597, 2584, 4181, 6765, 10946, 17711]
    fibtext = np.frombuffer(b"".join(bytes([65 + i]) * f for i, f in enumerate(fib)), np.uint8)
    lines = []
    for _ in range(3000):       # smoothed-quality look: '@' runs with a few original values, one line per read
        q = np.full(100, 64, np.uint8)
        for p in rng.integers(0, 100, rng.integers(0, 4)):
            q[p] = rng.integers(35, 75)
        lines.append(q.tobytes() + b"\n")
    t = {
        "smoothed": b"".join(lines),
        "lossless_like": rng.integers(35, 75, 150000).astype(np.uint8).tobytes(),
        "zeros": bytes(200000), "one": b"x", "two": b"ab", "five": b"aaaaa",
        "runs3": b"aaabbbcccdddd" * 5000,
        "random": rng.integers(0, 256, 130000).astype(np.uint8).tobytes(),      # stored blocks
        "run_edges": b"a" * (64512 * 2 + 7) + b"b" + b"a" * 259 + b"c" * 260 + b"d" * 261 + b"e" * 3 + b"f" * 2 + b"g" * 4,
        "fib": fibtext.tobytes(),                                                # Huffman depth > 15 before limiting
        "fib_shuffled": fibtext[rng.permutation(len(fibtext))].tobytes(),
        "pairs": np.repeat(rng.integers(33, 80, 40000), rng.integers(1, 6, 40000)).astype(np.uint8).tobytes(),
    }
    for n in (251, 252, 253, 504, 64511, 64512, 64513):
        t["len%d" % n] = rng.integers(60, 64, n).astype(np.uint8).tobytes()
    return t


@pytest.mark.parametrize("name", sorted(_texts()))
def test_stream_inflates_to_the_text(sim, tmp_path, name):
    text = _texts()[name]
    src, dst = str(tmp_path / "in.bin"), str(tmp_path / "out.z")
    open(src, "wb").write(text)
    subprocess.run([sim, src, dst], check=True, capture_output=True)
    z = open(dst, "rb").read()
    assert zlib.decompress(z) == text
    c = zlib.compressobj(9, zlib.DEFLATED, 15, 9, zlib.Z_RLE)
    ref = len(c.compress(text) + c.flush())
    nchunks = (len(text) + 64511) // 64512
    assert len(z) <= 1.02 * ref + 24 * nchunks + 16, (len(z), ref)
