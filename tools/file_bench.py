#!/usr/bin/env python3
"""File-level run of BASELINE configs[1] (5 Mbp genome, 100x, 100 bp => 4.6 M reads... here with Illumina-style read
names and qualities): `leon_b200 -file X -c` against the compiled reference `leon -file X -c` on the same file, both
timed by wall clock around the process, then the reference decodes OUR file and the reads / names are compared.

usage (GPU box): python tools/file_bench.py [--reads N] [--fastq] [--flags "-lossless"] [--runs 3] [--out gpurun_out/file_bench.json]
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CLI = os.path.join(ROOT, "leon_b200", "bin", "leon_b200")
REF = os.path.join(ROOT, "oracle", "_ref", "leon")


def write_file(path, bases, offs, quals):
    n = len(offs) - 1
    o = offs.astype(np.int64)
    x, y = 1200, 2000
    rng = np.random.Generator(np.random.PCG64(7))
    steps = rng.integers(1, 90, n)
    with open(path, "wb") as f:
        for lo in range(0, n, 1 << 16):
            hi = min(n, lo + (1 << 16))
            parts = []
            for i in range(lo, hi):
                y += int(steps[i])
                if y > 99000:
                    x, y = x + 1, 2000
                name = b"SRR1870605.%d HWI-ST1136:227:HS059ADXX:1:%d:%d:%d length=100" % (i + 1, 1101 + i // 200000, x, y)
                s = bases[o[i]:o[i + 1]].tobytes()
                if quals is None:
                    parts.append(b">%s\n%s\n" % (name, s))
                else:
                    parts.append(b"@%s\n%s\n+\n%s\n" % (name, s, quals[o[i]:o[i + 1]].tobytes()))
            f.write(b"".join(parts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=4_600_000)
    ap.add_argument("--fastq", action="store_true")
    ap.add_argument("--flags", default="")
    ap.add_argument("--runs", type=int, default=3)
    ap.add_argument("--inproc", type=int, default=0, help="also call leon_compress_file N times inside this process (CUDA context already up)")
    ap.add_argument("--no-reference", action="store_true")
    ap.add_argument("--no-decode", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "file_bench.json"))
    a = ap.parse_args()
    from leon_b200 import synth
    flags = a.flags.split()
    wd = tempfile.mkdtemp(prefix="filebench_")
    res = {"reads": a.reads, "fastq": a.fastq, "flags": flags, "cores": os.cpu_count()}
    try:
        glen = max(20000, a.reads * 100 // 92)      # ~92x, BASELINE configs[1] shape
        g = synth.make_genome(glen, 1)
        bases, offs, quals = synth.sample_reads(g, a.reads, 100, 2, n_rate=0.0002, with_quals=a.fastq)
        ext = "fastq" if a.fastq else "fasta"
        ours, ref = os.path.join(wd, "ours"), os.path.join(wd, "ref")
        os.makedirs(ours); os.makedirs(ref)
        inp = os.path.join(ours, "in." + ext)
        t = time.time()
        write_file(inp, bases, offs, quals)
        os.link(inp, os.path.join(ref, "in." + ext))
        res["input_bytes"] = os.path.getsize(inp)
        print("input written: %d bytes in %.1f s" % (res["input_bytes"], time.time() - t), flush=True)
        runs = []
        for r in range(a.runs):
            t = time.time()
            p = subprocess.run([CLI, "-file", inp, "-c"] + flags, capture_output=True, text=True)
            dt = time.time() - t
            assert p.returncode == 0, p.stdout + p.stderr
            runs.append(dt)
            print("ours run %d: %.3f s\n%s\n%s" % (r, dt, p.stdout.strip(), p.stderr.strip()), flush=True)
        res["ours_wall_s"] = runs
        if a.inproc:
            import leon_b200
            kw = dict(lossless="-lossless" in flags, noheader="-noheader" in flags, noqual="-noqual" in flags, seq_only="-seq-only" in flags)
            res["inproc"] = []
            for r in range(a.inproc):
                t = time.time()
                st = leon_b200.compress_file(inp, inp + ".inproc.leon", **kw)
                dt = time.time() - t
                res["inproc"].append({"wall_s": dt, "parse_ms": st["parse_ms"], "compress_ms": st["compress_ms"], "header_ms": st["header_ms"],
                                      "write_ms": st["write_ms"], "gpu_total_ms": st["gpu"]["total_ms"]})
                print("in-process call %d: %.3f s  (read %.0f ms, compress %.0f ms [GPU path %.0f ms, headers %.0f ms], write %.0f ms)" % (
                    r, dt, st["parse_ms"], st["compress_ms"], st["gpu"]["total_ms"], st["header_ms"], st["write_ms"]), flush=True)
            assert open(inp + ".inproc.leon", "rb").read() == open(inp + ".leon", "rb").read()
        res["ours_stdout"] = p.stdout.strip().splitlines()
        res["ours_output_bytes"] = os.path.getsize(inp + ".leon")
        # our own decompressor on our file: back to the input (names, reads; qualities when lossless)
        t = time.time()
        p = subprocess.run([CLI, "-file", inp + ".leon", "-d", "-out", os.path.join(ours, "ours.d")], capture_output=True, text=True)
        dt = time.time() - t
        assert p.returncode == 0, p.stdout + p.stderr
        res["ours_decode_wall_s"] = dt
        res["ours_decode_stdout"] = p.stdout.strip().splitlines()
        print("ours -d: %.3f s\n%s" % (dt, p.stdout.strip()), flush=True)
        if not ({"-seq-only", "-noheader", "-noqual"} & set(flags)):
            step = 4 if a.fastq else 2
            same = True
            with open(os.path.join(ours, "ours.d"), "rb") as fd, open(inp, "rb") as fi:
                i = 0
                for ld, li in zip(fd, fi):
                    if ld != li and (i % step != 3 or "-lossless" in flags):
                        same = False
                        break
                    i += 1
            res["ours_decode_equals_input"] = same and i == a.reads * step
            print("ours -d output equals the input (qualities compared only when lossless):", res["ours_decode_equals_input"], flush=True)
        if not a.no_reference and os.path.exists(REF):
            t = time.time()
            p = subprocess.run([REF, "-file", os.path.join(ref, "in." + ext), "-c", "-nb-cores", str(os.cpu_count())] + flags,
                               cwd=ref, capture_output=True, text=True)
            dt = time.time() - t
            assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
            res["reference_wall_s"] = dt
            res["reference_output_bytes"] = os.path.getsize(os.path.join(ref, "in.%s.leon" % ext))
            print("reference (-nb-cores %d): %.3f s, %d bytes" % (os.cpu_count(), dt, res["reference_output_bytes"]), flush=True)
            res["speedup_best"] = dt / min(runs)
        if not a.no_decode and os.path.exists(REF):
            t = time.time()
            p = subprocess.run([REF, "-file", inp + ".leon", "-d", "-nb-cores", str(os.cpu_count())], cwd=ours, capture_output=True, text=True)
            assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
            out_fastq = a.fastq and not ({"-seq-only", "-noqual"} & set(flags))
            dec = os.path.join(ours, "in.%s.d" % ("fastq" if out_fastq else "fasta"))
            step = 4 if out_fastq else 2
            ok_reads = ok_names = True
            if {"-seq-only", "-noheader"} & set(flags):
                res["decode_check"] = "skipped (header-less output)"
            else:
                o = offs.astype(np.int64)
                with open(dec, "rb") as fd, open(inp, "rb") as fi:
                    i = 0
                    for ld, li in zip(fd, fi):
                        k = i % step
                        if k == 0 and ld != li:
                            ok_names = False
                        if k == 1 and ld != li:
                            ok_reads = False
                        i += 1
                res["decode_check"] = {"records": i // step, "names_equal": ok_names, "reads_equal": ok_reads,
                                       "decode_wall_s": time.time() - t}
            print("reference decode of our file:", res["decode_check"], flush=True)
    finally:
        shutil.rmtree(wd, ignore_errors=True)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
