#!/usr/bin/env python3
"""Golden fixtures of the HEADER stream (and of whole files with headers), made by RUNNING THE REFERENCE.

Like make_golden.py (build container only; needs oracle/_ref built by oracle/build_ref.sh): for every case a FASTA/FASTQ
file with named reads is written, compressed with the reference `leon -c -nb-cores 1` WITHOUT -noheader, and the
leon/header datasets (firstheader, blocksizes, header_<id>) are pulled out with the reference's gatb-h5dump next to the
sequence-stream datasets, so that tests can (1) compare our host header coder byte for byte and (2) assemble a complete
container with our writer and have the reference decode it.  Output: tests/golden/hdrcase_<name>.npz.

usage: python tests/golden/make_golden_headers.py [--only NAME]
"""
import argparse
import os
import shutil
import subprocess
import sys
import tempfile
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as G  # noqa: E402

REF = os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle", "_ref")


def names_illumina(n, rng):
    x, y, out = 1200, 2000, []
    for i in range(n):
        y += int(rng.integers(1, 90))
        if y > 99000:
            x, y = x + int(rng.integers(1, 40)), 2000 + int(rng.integers(0, 50))
        tile = 1101 + (i // 700)
        out.append("HWI-ST1136:227:HS059ADXX:1:%d:%d:%d 1:N:0:ATCACG" % (tile, x, y))
    return out


def names_sra(n, rng):
    base = names_illumina(n, rng)
    return ["SRR1870605.%d %s length=100" % (i + 1, base[i].split(" ")[0]) for i in range(n)]


def names_zeros(n, rng):
    out = []
    for i in range(n):
        k = int(rng.integers(0, 4))
        if k == 0:
            out.append("read_%07d/1" % i)
        elif k == 1:
            out.append("read_%07d/2" % (i // 2))
        elif k == 2:
            out.append("read_000/%d" % (i % 3))
        else:
            out.append("read_%d_0000_x%03d" % (i * 7919 % 100000, i % 17))
    return out


def names_ragged(n, rng):
    """Field counts change from one header to the next, with numeric fields past the previous header's last field."""
    out = []
    for i in range(n):
        k = int(rng.integers(0, 6))
        if k == 0:
            out.append("r%d" % i)
        elif k == 1:
            out.append("r%d extra %d %d" % (i, i % 5, int(rng.integers(0, 3))))
        elif k == 2:
            out.append("r%d extra %d" % (i, i % 5))
        elif k == 3:
            out.append("%d" % int(rng.integers(0, 10 ** 12)))
        elif k == 4:
            out.append("r%d.%d.%d.%d 7 7" % (i, i % 2, 18446744073709551615 - i, i % 3))
        else:
            out.append("x y z %d 00 %d" % (int(rng.integers(0, 3)), int(rng.integers(0, 3))))
    return out


def names_long(n, rng):
    out = []
    for i in range(n):
        pad = "A" * int(rng.integers(0, 60))
        out.append("sample-%s|lane=%d|%s|idx=%d" % (pad, i % 4, "GATTACA" * int(rng.integers(1, 20)), i))
    return out


HCASES = {
    # name: (names generator, reads generator, extra leon flags, k, reads_per_block)
    "illumina_fa": (names_illumina, lambda: G.synth(31, 9000, 2400, 90), [], 31, 500),
    "sra_fq": (names_sra, lambda: G.synth(32, 9000, 2100, 100, fastq=True, n_rate=0.001), [], 31, 600),
    "zeros_fa": (names_zeros, lambda: G.synth(33, 6000, 1500, 80), [], 31, 400),
    "ragged_fa": (names_ragged, lambda: G.synth(34, 6000, 1800, 80), [], 31, 300),
    "long_fa": (names_long, lambda: G.synth(35, 6000, 900, 80), [], 31, 250),
    "one_block_fa": (names_sra, lambda: G.synth(36, 5000, 700, 80), [], 31, 50000),
}


def run_case(name, spec, leon, h5dump, outdir):
    namegen, readgen, flags, k, rpb = spec
    reads, quals = readgen()
    rng = np.random.Generator(np.random.PCG64(1000 + len(name)))
    names = namegen(len(reads), rng)
    scratch = tempfile.mkdtemp(prefix="goldenh_")
    try:
        ext = "fasta" if quals is None else "fastq"
        inp = os.path.join(scratch, "in." + ext)
        with open(inp, "wb") as f:
            for i, r in enumerate(reads):
                if quals is None:
                    f.write(b">" + names[i].encode() + b"\n" + r.tobytes() + b"\n")
                else:
                    f.write(b"@" + names[i].encode() + b"\n" + r.tobytes() + b"\n+\n" + quals[i].tobytes() + b"\n")
        cmd = [leon, "-file", inp, "-c", "-kmer-size", str(k), "-nb-cores", "1", "-reads", str(rpb)] + flags
        res = subprocess.run(cmd, cwd=scratch, capture_output=True, text=True)
        if res.returncode != 0 or "EXCEPTION" in res.stdout + res.stderr:
            raise RuntimeError("reference failed: %s\n%s\n%s" % (cmd, res.stdout[-2000:], res.stderr[-2000:]))
        lf = inp + ".leon"
        out = {}
        out["bloom"] = G.h5bin(h5dump, lf, "/leon/bloom", scratch)
        out["bloom_bits"] = np.uint64(int(G.h5attr(h5dump, lf, "/leon/bloom/size")))
        out["anchors_dict"] = G.h5bin(h5dump, lf, "/leon/anchors/anchorsDict", scratch)
        out["anchor_count"] = G.h5bin(h5dump, lf, "/leon/anchors/anchorAdress", scratch).view(np.uint32)[0]
        bs = G.h5bin(h5dump, lf, "/leon/dna/blocksizes", scratch).view(np.uint64)
        out["blocksizes"] = bs
        nblocks = len(bs) // 2
        dna = [G.h5bin(h5dump, lf, "/leon/dna/dna_%d" % b, scratch) for b in range(nblocks)]
        out["dna_concat"] = np.concatenate(dna)
        if quals is not None:
            q = [np.frombuffer(zlib.decompress(G.h5bin(h5dump, lf, "/leon/qual/qual_%d" % b, scratch).tobytes()), np.uint8)
                 for b in range(nblocks)]
            out["qual_text_sizes"] = np.array([len(x) for x in q], np.uint64)
            out["qual_text_concat"] = np.concatenate(q)
        for md in ("readcount", "totalDnaSize"):
            out["meta_" + md] = G.h5bin(h5dump, lf, "/metadata/" + md, scratch).view(np.uint64)[0]
        for md in ("minSequenceSize", "maxSequenceSize"):
            out["meta_" + md] = G.h5bin(h5dump, lf, "/metadata/" + md, scratch).view(np.int32)[0]
        out["infobyte"] = G.h5bin(h5dump, lf, "/metadata/infobyte", scratch)[0]
        # the header stream
        fhs = G.h5bin(h5dump, lf, "/leon/header/firstheadersize", scratch).view(np.uint64)[0]
        out["first_header"] = G.h5bin(h5dump, lf, "/leon/header/firstheader", scratch)
        assert len(out["first_header"]) == fhs
        hbs = G.h5bin(h5dump, lf, "/leon/header/blocksizes", scratch).view(np.uint64)
        hnb = G.h5bin(h5dump, lf, "/leon/header/nb_blocks", scratch).view(np.uint64)[0]
        assert hnb == len(hbs)
        out["header_blocksizes"] = hbs
        hb = [G.h5bin(h5dump, lf, "/leon/header/header_%d" % b, scratch) for b in range(len(hbs) // 2)]
        for b in range(len(hb)):
            assert len(hb[b]) == hbs[2 * b]
        out["header_concat"] = np.concatenate(hb)
        # Does the reference decode its own file back to the input, names included?  (Not always: with field counts that
        # change from header to header its decoder can throw, see leon_b200/csrc/headers.cpp on left-over state.)
        res = subprocess.run([leon, "-file", lf, "-d", "-nb-cores", "1"], cwd=scratch, capture_output=True, text=True)
        ok_names = ok_reads = False
        if res.returncode == 0:
            lines = [l.rstrip(b"\n") for l in open(os.path.join(scratch, "in.%s.d" % ext), "rb")]
            step = 2 if quals is None else 4
            ok_names = [l[1:] for l in lines[0::step]] == [n.encode() for n in names]
            ok_reads = lines[1::step] == [r.tobytes() for r in reads]
        out["ref_roundtrip_names_ok"] = np.bool_(ok_names)
        out["ref_roundtrip_reads_ok"] = np.bool_(ok_reads)

        offs = np.zeros(len(reads) + 1, np.uint64)
        offs[1:] = np.cumsum([len(r) for r in reads])
        hoffs = np.zeros(len(reads) + 1, np.uint64)
        hoffs[1:] = np.cumsum([len(n) for n in names])
        np.savez_compressed(
            os.path.join(outdir, "hdrcase_%s.npz" % name),
            bases=np.concatenate(reads), offsets=offs,
            quals=(np.concatenate(quals) if quals is not None else np.zeros(0, np.uint8)),
            headers=np.frombuffer("".join(names).encode(), np.uint8), header_offsets=hoffs,
            kmer_size=np.int32(k), reads_per_block=np.int32(rpb), flags=np.array(flags + ["-"]), **out)
        print("%-14s reads=%d header blocks=%d header bytes=%d -> %d  reference decodes its own file: names %s reads %s" % (
            name, len(reads), len(hb), int(hoffs[-1]), len(out["header_concat"]), ok_names, ok_reads))
    finally:
        shutil.rmtree(scratch, ignore_errors=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--leon", default=os.path.join(REF, "leon"))
    ap.add_argument("--h5dump", default=os.path.join(REF, "gatb-h5dump"))
    ap.add_argument("--only", default=None)
    a = ap.parse_args()
    for name, spec in HCASES.items():
        if a.only and a.only != name:
            continue
        run_case(name, spec, a.leon, a.h5dump, HERE)


if __name__ == "__main__":
    main()
