#!/usr/bin/env python3
"""One run of the hot path on a BASELINE-shaped synthetic set of arbitrary size, with size-independent checks (no
oracle at this size): block table sums, read counts, anchors found, identical output on a second run.

    python tools/run_scale.py --genome 100000000 --reads 33333333 --len 150 [--fastq] [--k 31]
"""
import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome", type=int, default=100_000_000)
    ap.add_argument("--reads", type=int, default=33_333_333)
    ap.add_argument("--len", type=int, default=150)
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--fastq", action="store_true")
    ap.add_argument("--runs", type=int, default=2)
    a = ap.parse_args()
    import leon_b200
    from leon_b200 import synth
    t0 = time.time()
    g = synth.make_genome(a.genome, 3)
    pin_b = leon_b200.api.PinnedArray(a.reads * a.len)
    pin_q = leon_b200.api.PinnedArray(a.reads * a.len) if a.fastq else None
    bases, offs, quals = synth.sample_reads(g, a.reads, a.len, 4, n_rate=0.001, out_bases=pin_b.array, with_quals=a.fastq,
                                            out_quals=pin_q.array if a.fastq else None, chunk=1 << 20)
    print("generated %d reads in %.1f s" % (a.reads, time.time() - t0), flush=True)
    eng = leon_b200.Leon(kmer_size=a.k, reads_per_block=50000, has_qual=a.fastq)
    digests = []
    for i in range(a.runs):
        t1 = time.time()
        r = eng.compress(bases, offs, quals, copy=False)
        wall = time.time() - t1
        tm = eng.timings()
        kt = eng.kernel_times()
        h = hashlib.sha256()
        for key in ("bloom", "anchors_dict", "dna", "blocksizes", "anchor_pos"):
            h.update(np.ascontiguousarray(r[key]).tobytes())
        digests.append(h.hexdigest())
        bs = r["blocksizes"]
        assert int(bs[0::2].sum()) == len(r["dna"]), "block sizes do not add up to the payload"
        assert int(bs[1::2].sum()) == a.reads == r["read_count"], "block read counts do not add up"
        assert r["nblocks"] == (a.reads + 49999) // 50000
        assert r["anchor_count"] > 0 and (r["anchor_pos"] >= 0).mean() > 0.9
        assert r["total_dna_size"] == a.reads * a.len
        out = {"run": i, "wall_s": wall, "reads_per_s": a.reads / wall, "bloom_bytes": len(r["bloom"]), "anchors": r["anchor_count"],
               "dna_bytes": len(r["dna"]), "bytes_per_read": len(r["dna"]) / a.reads, "cutoff": r["cutoff"],
               "solid": r["nb_solids_cutoff"], "phases_ms": {k: round(v, 2) for k, v in tm.items() if k.endswith("_ms")},
               "kernels_ms": {k: round(v[0], 2) for k, v in sorted(kt.items(), key=lambda kv: -kv[1][0])}}
        print(json.dumps(out), flush=True)
    assert len(set(digests)) == 1, "outputs of two runs differ"
    print("all property checks passed; digest", digests[0][:16])
    eng.close()


if __name__ == "__main__":
    main()
