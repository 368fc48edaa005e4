#!/usr/bin/env python3
"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE.

Runs only in the build container (needs /root/reference built, see
oracle/build_ref.sh).  For every case it writes a FASTA/FASTQ file to a scratch
directory, runs the reference `leon -c ... -nb-cores 1` (deterministic, SURVEY.md
section 8c), pulls the payload datasets out of the `.leon` HDF5 container with the
reference's own `gatb-h5dump -b`, inflates the quality blocks, and stores inputs +
expected outputs in one compressed `case_<name>.npz`.

Nothing here is imported by the product; tests only read the .npz files.

usage: python tests/golden/make_golden.py [--leon BIN] [--h5dump BIN] [--only NAME]
"""
import argparse
import os
import re
import shutil
import subprocess
import sys
import tempfile
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = {65: 84, 67: 71, 71: 67, 84: 65, 78: 78}


def revcomp(a):
    lut = np.zeros(256, np.uint8)
    for k, v in COMP.items():
        lut[k] = v
    return lut[a[::-1]]


def synth(seed, genome_len, n_reads, read_len, sub=0.005, n_rate=0.0, n_runs=0.0,
          var_len=False, short_frac=0.0, k=31, fastq=False, indel=0.0):
    """Reads sampled with errors from a uniform random genome (both strands)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    genome = ACGT[rng.integers(0, 4, genome_len)]
    reads, quals = [], []
    for _ in range(n_reads):
        L = read_len
        if var_len:
            L = int(rng.integers(k, read_len + 1))
        if short_frac and rng.random() < short_frac:
            L = int(rng.integers(1, k))
        p = int(rng.integers(0, genome_len - L + 1))
        r = genome[p:p + L].copy()
        if rng.random() < 0.5:
            r = revcomp(r)
        q = rng.choice(np.array([30, 32, 35, 37, 38, 40, 40, 40]), L)
        tail = max(0, L - 20)
        q[tail:] = np.maximum(2, q[tail:] - rng.integers(0, 11, L - tail))
        err = rng.random(L) < sub
        for j in np.nonzero(err)[0]:
            r[j] = ACGT[(np.searchsorted(ACGT, r[j]) + rng.integers(1, 4)) % 4]
            q[j] = rng.integers(2, 16)
        if indel and rng.random() < indel * L:
            j = int(rng.integers(1, L - 1))
            r = np.concatenate([r[:j], r[j + 1:], ACGT[rng.integers(0, 4, 1)]])
        if n_rate:
            nm = rng.random(L) < n_rate
            r[nm] = 78
            q[nm] = 2
        if n_runs and rng.random() < n_runs and L > 12:
            j = int(rng.integers(0, L - 5))
            e = min(L, j + int(rng.integers(5, 31)))
            r[j:e] = 78
            q[j:e] = 2
        reads.append(r)
        quals.append((q + 33).astype(np.uint8))
    return reads, (quals if fastq else None)


def toy_reads():
    reads, cur = [], []
    with open("/root/reference/data/toy.fasta", "rb") as f:
        for line in f:
            line = line.strip()
            if line.startswith(b">"):
                if cur:
                    reads.append(np.frombuffer(b"".join(cur), np.uint8))
                cur = []
            elif line:
                cur.append(line)
    if cur:
        reads.append(np.frombuffer(b"".join(cur), np.uint8))
    return reads, None


CASES = {
    # name: (generator, leon flags, k, reads_per_block)
    "toy_k31": (toy_reads, ["-seq-only"], 31, 50000),
    "fa_k31_blocks": (lambda: synth(11, 20000, 5000, 100), ["-seq-only"], 31, 1000),
    "fa_k31_abund": (lambda: synth(12, 12000, 3000, 100), ["-seq-only", "-abundance", "3"], 31, 1000),
    "fa_k21_var": (lambda: synth(13, 15000, 3000, 120, var_len=True, k=21, n_rate=0.002), ["-seq-only"], 21, 700),
    "fa_k31_edge": (lambda: synth(14, 8000, 2500, 90, n_rate=0.003, n_runs=0.02, short_frac=0.03, var_len=True,
                                  indel=0.0005), ["-seq-only"], 31, 600),
    "fq_k31_lossy": (lambda: synth(15, 10000, 3000, 100, n_rate=0.001, n_runs=0.01, short_frac=0.01, fastq=True),
                     ["-noheader"], 31, 1000),
    "fq_k31_lossless": (lambda: synth(16, 8000, 2000, 100, n_rate=0.001, fastq=True),
                        ["-noheader", "-lossless"], 31, 1000),
    "fa_k31_lowcov": (lambda: synth(17, 200000, 2000, 100), ["-seq-only"], 31, 50000),
}
CASES64 = {
    "fa_k63": (lambda: synth(21, 15000, 3000, 150, k=63, n_rate=0.001), ["-seq-only"], 63, 1000),
    "fq_k47_lossy": (lambda: synth(22, 8000, 2000, 120, k=47, fastq=True, n_rate=0.001), ["-noheader"], 47, 800),
}


def write_input(path, reads, quals):
    with open(path, "wb") as f:
        for i, r in enumerate(reads):
            if quals is None:
                f.write(b">r%d\n" % i + r.tobytes() + b"\n")
            else:
                f.write(b"@r%d\n" % i + r.tobytes() + b"\n+\n" + quals[i].tobytes() + b"\n")


def h5bin(h5dump, leon_file, dset, scratch):
    out = os.path.join(scratch, "dump.bin")
    if os.path.exists(out):
        os.remove(out)
    subprocess.run([h5dump, "-d", dset, "-b", "LE", "-o", out, leon_file], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return np.fromfile(out, dtype=np.uint8)


def h5attr(h5dump, leon_file, attr):
    txt = subprocess.run([h5dump, "-a", attr, leon_file], check=True, capture_output=True, text=True).stdout
    return re.search(r'\(0\): "(.*)"', txt).group(1)


def run_case(name, spec, leon, h5dump, outdir):
    gen, flags, k, rpb = spec
    reads, quals = gen()
    scratch = tempfile.mkdtemp(prefix="golden_")
    try:
        ext = "fasta" if quals is None else "fastq"
        inp = os.path.join(scratch, "in." + ext)
        write_input(inp, reads, quals)
        cmd = [leon, "-file", inp, "-c", "-kmer-size", str(k), "-nb-cores", "1", "-reads", str(rpb)] + flags
        res = subprocess.run(cmd, cwd=scratch, capture_output=True, text=True)
        if res.returncode != 0 or "EXCEPTION" in res.stdout + res.stderr:
            raise RuntimeError("reference failed: %s\n%s\n%s" % (cmd, res.stdout[-2000:], res.stderr[-2000:]))
        log = res.stdout
        m = re.search(r"cut-off(?: \(auto\))?\s*:\s*(\d+)", log)
        cutoff = int(m.group(1))
        auto = "cut-off (auto)" in log
        nbsolid = int(re.search(r"nb solid kmers\s*:\s*(\d+)", log).group(1))
        lf = inp + ".leon"
        out = {}
        out["bloom"] = h5bin(h5dump, lf, "/leon/bloom", scratch)
        out["bloom_bits"] = np.uint64(int(h5attr(h5dump, lf, "/leon/bloom/size")))
        out["anchors_dict"] = h5bin(h5dump, lf, "/leon/anchors/anchorsDict", scratch)
        out["anchor_count"] = h5bin(h5dump, lf, "/leon/anchors/anchorAdress", scratch).view(np.uint32)[0]
        bs = h5bin(h5dump, lf, "/leon/dna/blocksizes", scratch).view(np.uint64)
        out["blocksizes"] = bs
        nblocks = len(bs) // 2
        dna = [h5bin(h5dump, lf, "/leon/dna/dna_%d" % b, scratch) for b in range(nblocks)]
        for b in range(nblocks):
            assert len(dna[b]) == bs[2 * b], (name, b, len(dna[b]), bs[2 * b])
        out["dna_concat"] = np.concatenate(dna) if dna else np.zeros(0, np.uint8)
        if quals is not None:
            q = []
            for b in range(nblocks):
                z = h5bin(h5dump, lf, "/leon/qual/qual_%d" % b, scratch)
                q.append(np.frombuffer(zlib.decompress(z.tobytes()), np.uint8))
            out["qual_text_sizes"] = np.array([len(x) for x in q], np.uint64)
            out["qual_text_concat"] = np.concatenate(q)
        for md in ("readcount", "totalDnaSize"):
            out["meta_" + md] = h5bin(h5dump, lf, "/metadata/" + md, scratch).view(np.uint64)[0]
        for md in ("minSequenceSize", "maxSequenceSize"):
            out["meta_" + md] = h5bin(h5dump, lf, "/metadata/" + md, scratch).view(np.int32)[0]
        # decompress with the reference as a sanity check of the fixture itself
        res = subprocess.run([leon, "-file", lf, "-d", "-nb-cores", "1"], cwd=scratch, capture_output=True, text=True)
        assert res.returncode == 0, res.stdout + res.stderr
        dec = os.path.join(scratch, "in.%s.d" % ext)
        # quirk: in header-less mode the reference decoder prints one stray "> <n>" / "@ <n>" line at the
        # end of every full block, so drop the numbered pseudo-headers instead of counting lines
        hdr = re.compile(rb"^[>@] \d+$")
        lines = [l.rstrip(b"\n") for l in open(dec, "rb")]
        lines = [l for l in lines if not hdr.match(l)]
        got = lines if quals is None else lines[0::3]
        assert got == [r.tobytes() for r in reads], "reference round trip of the fixture failed"
        if quals is not None and "-lossless" in flags:
            assert lines[2::3] == [x.tobytes() for x in quals]

        offs = np.zeros(len(reads) + 1, np.uint64)
        offs[1:] = np.cumsum([len(r) for r in reads])
        np.savez_compressed(
            os.path.join(outdir, "case_%s.npz" % name),
            bases=np.concatenate(reads), offsets=offs,
            quals=(np.concatenate(quals) if quals is not None else np.zeros(0, np.uint8)),
            kmer_size=np.int32(k), reads_per_block=np.int32(rpb), flags=np.array(flags),
            cutoff=np.int32(cutoff), auto_cutoff=np.bool_(auto), nb_solid=np.uint64(nbsolid), **out)
        print("%-18s reads=%d blocks=%d cutoff=%d%s nbsolid=%d bloom=%dB anchors=%d dict=%dB dna=%dB" % (
            name, len(reads), nblocks, cutoff, "(auto)" if auto else "", nbsolid, len(out["bloom"]),
            out["anchor_count"], len(out["anchors_dict"]), len(out["dna_concat"])))
    finally:
        shutil.rmtree(scratch, ignore_errors=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--leon", default="/tmp/leon_build/bin/leon")
    ap.add_argument("--leon64", default="/tmp/leon_build64/bin/leon")
    ap.add_argument("--h5dump", default="/tmp/leon_build/ext/gatb-core/bin/Release/gatb-h5dump")
    ap.add_argument("--only", default=None)
    a = ap.parse_args()
    for name, spec in CASES.items():
        if a.only and a.only != name:
            continue
        run_case(name, spec, a.leon, a.h5dump, HERE)
    if os.path.exists(a.leon64):
        for name, spec in CASES64.items():
            if a.only and a.only != name:
                continue
            run_case(name, spec, a.leon64, a.h5dump, HERE)
    else:
        print("k>31 reference build not found at %s: CASES64 skipped" % a.leon64, file=sys.stderr)


if __name__ == "__main__":
    main()
