#!/usr/bin/env python3
"""TEST INFRASTRUCTURE.  Full-size byte parity: writes the seeded BASELINE workload to a file, runs the UNMODIFIED reference
(oracle/_ref/leon -c -nb-cores 1: deterministic) on it and stores the digest of its sequence (+ quality)
streams in tests/golden/digest_<name>.json.  bench.py recomputes the same digest from the GPU output at every N
(`output_digest`, `digest_ok`).  Runs in the build container only (minutes of one host core per workload).

    python tools/make_digest.py configs[2] [configs[1] ...]
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from leon_b200 import workloads as W  # noqa: E402


def main():
    names = sys.argv[1:] or ["configs[2]"]
    ref = os.path.join(ROOT, "oracle", "_ref", "leon")
    for name in names:
        c = W.CONFIGS[name]
        tmp = tempfile.mkdtemp(prefix="leondig_", dir=os.environ.get("LEON_TMP", "/tmp"))
        try:
            t0 = time.time()
            bases, offs, quals = W.generate(name)
            path = os.path.join(tmp, "w.fastq" if c["fastq"] else "w.fasta")
            W.write_fixed(path, bases, c["read_len"], quals)
            t1 = time.time()
            cmd = [ref, "-file", path, "-c", "-kmer-size", str(W.KMER), "-nb-cores", "1", "-verbose", "0"] + c["ref_flags"]
            r = subprocess.run(cmd, cwd=tmp, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                raise SystemExit("reference failed: " + r.stdout[-500:])
            t2 = time.time()
            d = W.digest_of_container(path + ".leon", c["fastq"])
            d.update(workload=c["text"], reads=c["reads"], command=" ".join(["leon"] + cmd[3:]),
                     reference_wall_s=round(t2 - t1, 1), made_by="tools/make_digest.py")
            out = os.path.join(ROOT, "tests", "golden", "digest_%s.json" % name.replace("[", "").replace("]", ""))
            json.dump(d, open(out, "w"), indent=1)
            print(name, "generated %.0f s, reference %.0f s ->" % (t1 - t0, t2 - t1), out, flush=True)
        finally:
            shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
