#!/usr/bin/env python3
"""TEST INFRASTRUCTURE.  Digest of the reference's output (oracle/_ref/leon -nb-cores 1, deterministic) for the configs[3]
workload at a given scale (default: the full 33.3 M reads; about 10 GB of FASTQ and tens of minutes of one host core).
Build container only.     python tools/make_digest_c3.py [reads] [genome]"""
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from leon_b200 import synth, workloads as W  # noqa: E402


def main():
    c = W.CONFIGS["configs[3]"]
    reads = int(sys.argv[1]) if len(sys.argv) > 1 else c["reads"]
    genome = int(sys.argv[2]) if len(sys.argv) > 2 else c["genome"]
    ref = os.path.join(ROOT, "oracle", "_ref", "leon")
    tmp = tempfile.mkdtemp(prefix="leondig3_", dir=os.environ.get("LEON_TMP", "/tmp"))
    try:
        t0 = time.time()
        g = synth.make_genome(genome, c["gseed"])
        path = os.path.join(tmp, "w.fastq")
        nchunks = (reads + W.C3_CHUNK - 1) // W.C3_CHUNK
        for ch in range(nchunks):
            b, lens, q = W.c3_chunk(g, ch, reads, c["read_len"], c["rseed"], True)
            import numpy as np
            offs = np.zeros(len(lens) + 1, np.uint64)
            np.cumsum(lens, out=offs[1:])
            W.write_ragged(path, b, offs, q, first_index=ch * W.C3_CHUNK, append=ch > 0)
        t1 = time.time()
        print("file written in %.0f s" % (t1 - t0), flush=True)
        cmd = [ref, "-file", path, "-c", "-kmer-size", str(W.KMER), "-nb-cores", "1", "-verbose", "0"] + c["ref_flags"]
        r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise SystemExit("reference failed: " + r.stdout[-500:])
        t2 = time.time()
        d = W.digest_of_container(path + ".leon", True)
        d.update(workload=c["text"], reads=reads, genome=genome, command=" ".join(["leon"] + cmd[3:]), reference_wall_s=round(t2 - t1, 1),
                 made_by="tools/make_digest_c3.py")
        name = "digest_configs3.json" if reads == c["reads"] and genome == c["genome"] else "digest_configs3_%d.json" % reads
        json.dump(d, open(os.path.join(ROOT, "tests", "golden", name), "w"), indent=1)
        print("reference %.0f s ->" % (t2 - t1), name, flush=True)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
