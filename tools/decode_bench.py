"""Sequence-stream decoding at configs[1] scale: the host-thread decoder (leon_decode_dna, all cores) beside the GPU
decoder (leon_decode_dna_device, one warp per block).  Both are checked against the input reads.
  python tools/decode_bench.py [--reads 4600000] [--read-len 100] [--genome 4600000] [--rpb 50000] [--out FILE]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=4600000)
    ap.add_argument("--read-len", type=int, default=100)
    ap.add_argument("--genome", type=int, default=4600000)
    ap.add_argument("--rpb", type=int, default=50000)
    ap.add_argument("--kmer-size", type=int, default=31)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import leon_b200
    from leon_b200 import synth
    g = synth.make_genome(a.genome, 1)
    bases, offs, _ = synth.sample_reads(g, a.reads, a.read_len, 2)
    eng = leon_b200.Leon(kmer_size=a.kmer_size, reads_per_block=a.rpb)
    r = eng.compress(bases, offs)
    eng.close()
    text = bases.reshape(a.reads, a.read_len)
    res = dict(reads=a.reads, read_len=a.read_len, blocks=int(r["nblocks"]), dna_bytes=int(len(r["dna"])), cores=os.cpu_count())

    def same(lines):
        if len(lines) != a.reads:
            return False
        got = np.frombuffer(b"".join(lines), np.uint8)
        return got.size == text.size and bool(np.array_equal(got.reshape(text.shape), text))

    for rep in range(2):
        t = time.perf_counter()
        lines, ms = leon_b200.decode_dna_device(r, a.kmer_size, with_time=True)
        wall = time.perf_counter() - t
        res["device_call_s_run%d" % rep] = wall
        res["device_kernel_ms_run%d" % rep] = ms
    res["device_equals_input"] = same(lines)
    res["device_reads_per_s_kernel"] = a.reads / (ms / 1e3)
    t = time.perf_counter()
    lines = leon_b200.decode_dna(r, a.kmer_size, 0)
    res["host_call_s"] = time.perf_counter() - t
    res["host_equals_input"] = same(lines)
    res["note"] = "call_s includes turning the decoded text into Python lines (same for both); kernel_ms is the device time of k_decode_blocks"
    line = json.dumps(res)
    print(line)
    if a.out:
        open(a.out, "w").write(line + "\n")


if __name__ == "__main__":
    main()
