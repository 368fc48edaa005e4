"""File level on the GPU box: `leon_b200 -file X -c` (the command line tool over leon_compress_file) against the compiled
reference on the same file: (1) every dataset and attribute of our `.leon` equals the one in the container the
reference writes with -nb-cores 1 (quality blocks compared after inflating), (2) the reference decodes our file back to
the input names, reads and (lossless) qualities."""
import os
import subprocess
import zlib

import numpy as np
import pytest

from container_util import H5Mini, ref_attr, ref_dataset, ref_h5dump, ref_leon, REF_DIR, ROOT

pytestmark = pytest.mark.gpu
CLI = os.path.join(ROOT, "leon_b200", "bin", "leon_b200")


def _names(n, rng):
    x, y, out = 1200, 2000, []
    for i in range(n):
        y += int(rng.integers(1, 90))
        if y > 99000:
            x, y = x + int(rng.integers(1, 40)), 2000
        out.append("SRR1870605.%d HWI-ST1136:227:HS059ADXX:1:%d:%d:%d length=100" % (i + 1, 1101 + i // 7000, x, y))
    return out


def _write_file(path, fastq, nreads, seed, line_width=0):
    from leon_b200 import synth
    rng = np.random.Generator(np.random.PCG64(seed))
    g = synth.make_genome(60000, seed)
    bases, offs, quals = synth.sample_reads(g, nreads, 100, seed + 1, n_rate=0.0005, with_quals=fastq)
    names = _names(nreads, rng)
    o = offs.astype(np.int64)
    reads = [bases[o[i]:o[i + 1]].tobytes() for i in range(nreads)]
    qs = [quals[o[i]:o[i + 1]].tobytes() for i in range(nreads)] if fastq else None
    with open(path, "wb") as f:
        for i in range(nreads):
            if fastq:
                f.write(b"@" + names[i].encode() + b"\n" + reads[i] + b"\n+\n" + qs[i] + b"\n")
            elif line_width:
                r = reads[i]
                f.write(b">" + names[i].encode() + b"\n" + b"\n".join(r[j:j + line_width] for j in range(0, len(r), line_width)) + b"\n")
            else:
                f.write(b">" + names[i].encode() + b"\n" + reads[i] + b"\n")
    return [n.encode() for n in names], reads, qs


def _decode(leon_file, out_fastq, wd):
    res = subprocess.run([ref_leon(31), "-file", leon_file, "-d", "-nb-cores", "1"], cwd=wd, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-1500:] + res.stderr[-1500:]
    # the decoder names its output after the stored type: <prefix>.fastq.d or <prefix>.fasta.d (C/Leon.cpp:1437-1448)
    return [l.rstrip(b"\n") for l in open(os.path.join(wd, "in.%s.d" % ("fastq" if out_fastq else "fasta")), "rb")]


@pytest.mark.parametrize("ext,flags", [("fastq", ["-lossless"]), ("fastq", []), ("fasta", []), ("fastq", ["-seq-only"]),
                                       ("fastq", ["-noheader"]), ("fastq", ["-noqual"])])
def test_cli_file_equals_reference_container(tmp_path, ext, flags):
    if ref_leon(31) is None or ref_h5dump() is None:
        pytest.skip("oracle/_ref not built")
    fastq = ext == "fastq"
    ours_dir, ref_dir = tmp_path / "ours", tmp_path / "ref"
    ours_dir.mkdir(); ref_dir.mkdir()
    names, reads, quals = _write_file(str(ours_dir / ("in." + ext)), fastq, 23000, 5, line_width=0 if fastq else 60)
    os.link(str(ours_dir / ("in." + ext)), str(ref_dir / ("in." + ext)))
    common = ["-c", "-kmer-size", "31", "-reads", "5000"] + flags
    res = subprocess.run([CLI, "-file", str(ours_dir / ("in." + ext))] + common, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout + res.stderr
    res = subprocess.run([ref_leon(31), "-file", str(ref_dir / ("in." + ext)), "-nb-cores", "1"] + common, cwd=str(ref_dir),
                         capture_output=True, text=True, timeout=900)
    assert res.returncode == 0 and "EXCEPTION" not in res.stdout + res.stderr, res.stdout[-1500:] + res.stderr[-1500:]
    mine, theirs = str(ours_dir / ("in.%s.leon" % ext)), str(ref_dir / ("in.%s.leon" % ext))

    h = H5Mini(mine).root
    compared = 0

    def walk(node, path):
        nonlocal compared
        for name, child in node.items():
            if name.startswith("_"):
                continue
            if "attrs" in child:        # dataset
                d = ref_dataset(theirs, path + "/" + name, str(tmp_path))
                if name.startswith("qual_"):
                    assert zlib.decompress(child["data"].tobytes()) == zlib.decompress(d.tobytes()), path + "/" + name
                else:
                    assert np.array_equal(child["data"], d), path + "/" + name
                for a, v in child["attrs"].items():
                    if not (name.startswith("qual_") and a == "size"):
                        assert ref_attr(theirs, path + "/" + name + "/" + a) == v, (path, name, a)
                compared += 1
            else:
                walk(child, path + "/" + name)

    walk(h, "")
    nblocks = 5
    seq_only, noheader, noqual = "-seq-only" in flags, "-noheader" in flags or "-seq-only" in flags, "-noqual" in flags or "-seq-only" in flags
    expect = 6 + 1 + 3 + (nblocks + 2) + (0 if noheader else nblocks + 4) + (nblocks if fastq and not noqual else 0)
    assert compared == expect, (compared, expect)

    out_fastq = fastq and not noqual
    lines = _decode(mine, out_fastq, str(ours_dir))
    if noheader:
        import re
        hdr = re.compile(rb"^[>@] \d+$")
        lines = [l for l in lines if not hdr.match(l)]
        got_reads = lines[0::3] if out_fastq else lines
    else:
        step = 4 if out_fastq else 2
        assert [l[1:] for l in lines[0::step]] == names
        got_reads = lines[1::step]
        if "-lossless" in flags:
            assert lines[3::4] == quals
    assert got_reads == reads


@pytest.mark.parametrize("ext,flags", [("fastq", []), ("fastq", ["-lossless"]), ("fasta", [])])
def test_cli_streaming_writes_the_same_container(tmp_path, ext, flags):
    """`-stream-reads N -count-passes P`: the file fed to the GPU in batches of N reads (streaming entry points) gives the same
    `.leon` file, byte for byte, as the whole set at once (which equals the reference's, test above)."""
    fastq = ext == "fastq"
    a, b = tmp_path / "a", tmp_path / "b"
    a.mkdir(); b.mkdir()
    _write_file(str(a / ("in." + ext)), fastq, 23000, 9)
    os.link(str(a / ("in." + ext)), str(b / ("in." + ext)))
    common = ["-c", "-kmer-size", "31", "-reads", "2000"] + flags
    r1 = subprocess.run([CLI, "-file", str(a / ("in." + ext))] + common, capture_output=True, text=True, timeout=600)
    r2 = subprocess.run([CLI, "-file", str(b / ("in." + ext))] + common + ["-stream-reads", "7000", "-count-passes", "3"],
                        capture_output=True, text=True, timeout=600)
    assert r1.returncode == 0 and r2.returncode == 0, r1.stderr + r2.stderr
    assert (a / ("in.%s.leon" % ext)).read_bytes() == (b / ("in.%s.leon" % ext)).read_bytes()


def test_cli_gzip_input_and_output_name(tmp_path):
    """`leon -file x.fastq.gz -c` (the reference reads through gzopen): same streams as the plain file, and the output is
    <dir>/x.fastq.leon (C/Leon.cpp:466-479)."""
    import gzip
    import leon_b200
    d = tmp_path / "g"
    d.mkdir()
    _write_file(str(d / "in.fastq"), True, 6000, 11)
    (d / "x.fastq.gz").write_bytes(gzip.compress((d / "in.fastq").read_bytes()))
    common = ["-c", "-kmer-size", "31", "-reads", "2000"]
    r1 = subprocess.run([CLI, "-file", str(d / "in.fastq")] + common, capture_output=True, text=True, timeout=600)
    r2 = subprocess.run([CLI, "-file", str(d / "x.fastq.gz")] + common, capture_output=True, text=True, timeout=600)
    assert r1.returncode == 0 and r2.returncode == 0, r1.stderr + r2.stderr
    assert (d / "x.fastq.leon").exists() and not (d / "x.fastq.gz.leon").exists()
    with leon_b200.Container(str(d / "in.fastq.leon")) as c1, leon_b200.Container(str(d / "x.fastq.leon")) as c2:
        for ds in ("leon/bloom", "leon/anchors/anchorsDict", "leon/dna/blocksizes", "leon/dna/dna_0", "leon/dna/dna_2", "leon/qual/qual_1",
                   "leon/header/header_0"):
            assert np.array_equal(c1.dataset(ds), c2.dataset(ds)), ds


def test_cli_compress_summary_has_the_reference_keys(tmp_path):
    """The lines users and CI logs read (C/Leon.cpp:402-416, 726-767, 970-1015) are printed by `leon -c`, the values of the
    counting stage equal the reference's."""
    _write_file(str(tmp_path / "in.fastq"), True, 12000, 13)
    leon = os.path.join(ROOT, "leon_b200", "bin", "leon")
    r = subprocess.run([leon, "-file", str(tmp_path / "in.fastq"), "-c", "-kmer-size", "31"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    for key in ("Input format: FastQ", "Quality compression: lossy mode", "End Header compression", "cut-off (auto)", "nb solid kmers",
                "End Sequence compression", "# blocks", "# sequences", "# nucleotides", "Reads per anchor", "Read without anchor",
                "End compression", "Time:", "Speed:", "overall", "header only", "Sequence only", "Quality only"):
        assert key in r.stdout, key
    if ref_leon(31) is None:
        pytest.skip("oracle/_ref not built")
    d = tmp_path / "ref"
    d.mkdir()
    os.link(str(tmp_path / "in.fastq"), str(d / "in.fastq"))
    rr = subprocess.run([ref_leon(31), "-file", str(d / "in.fastq"), "-c", "-kmer-size", "31", "-nb-cores", "1"], cwd=str(d),
                        capture_output=True, text=True, timeout=900)
    ref_out = rr.stdout.replace("\r", "\n")

    def value(text, key):
        for ln in text.split("\n"):
            if ln.strip().startswith(key):
                return ln.split(":")[-1].strip()
        return None
    for key in ("cut-off (auto)", "nb solid kmers", "# sequences", "# nucleotides"):
        assert value(r.stdout, key) == value(ref_out, key), key
