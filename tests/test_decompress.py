"""Decompression side on CPU (leon_decode_dna, leon_headers_decode, leon_decompress_file).
(1) The sequence and header streams the REFERENCE wrote (golden fixtures) decode to the input reads and names.
(2) leon_decompress_file on containers written by OUR writer and by the REFERENCE produces byte for byte the file the
    reference's own `leon -d -nb-cores 1` writes (header-less numbering quirks included)."""
import os
import subprocess

import numpy as np
import pytest

from container_util import reads_of, ref_leon, result_from_golden
from golden_util import case_names, load_case
from test_headers import hdr_cases, load_hdr

CLI = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "leon_b200", "bin", "leon_b200")


@pytest.mark.parametrize("nthreads", [1, 3])
@pytest.mark.parametrize("name", case_names())
def test_reference_sequence_streams_decode_to_the_reads(name, nthreads):
    import leon_b200
    c = load_case(name)
    r = result_from_golden(c)
    assert leon_b200.decode_dna(r, c["k"], nthreads) == reads_of(c["bases"], c["offsets"])


@pytest.mark.parametrize("name", hdr_cases())
def test_reference_header_streams_decode_to_the_names(name):
    import leon_b200
    c = load_hdr(name)
    h = dict(first_header=c["first_header"].tobytes(), blocksizes=c["header_blocksizes"], data=c["header_concat"])
    if not bool(c["ref_roundtrip_names_ok"]):
        with pytest.raises(leon_b200.LeonError):      # the reference's decoder throws on this stream, ours reports it
            leon_b200.decode_headers(h, 2)
        return
    for nt in (1, 4):
        assert leon_b200.decode_headers(h, nt) == reads_of(c["headers"], c["header_offsets"])


def test_corrupt_streams_are_errors_not_crashes():
    import leon_b200
    c = load_case("fa_k31_blocks")
    r = result_from_golden(c)
    rng = np.random.default_rng(5)
    for trial in range(20):
        bad = dict(r)
        dna = r["dna"].copy()
        pos = rng.integers(0, len(dna), 40)
        dna[pos] = rng.integers(0, 256, 40)
        bad["dna"] = dna
        bad["max_read_size"] = int(c["meta_maxSequenceSize"])
        try:
            leon_b200.decode_dna(bad, c["k"], 2)       # garbage in: either an error or some reads, never a crash
        except leon_b200.LeonError:
            pass


def _ours_vs_reference(tmp_path, leon_file, k):
    import leon_b200
    ours = tmp_path / "ours.out"
    st = leon_b200.decompress_file(leon_file, str(ours))
    res = subprocess.run([ref_leon(k), "-file", leon_file, "-d", "-nb-cores", "1"], cwd=os.path.dirname(leon_file),
                         capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-1500:] + res.stderr[-1500:]
    theirs = [f for f in os.listdir(os.path.dirname(leon_file)) if f.endswith(".d")]
    assert len(theirs) == 1
    assert open(ours, "rb").read() == open(os.path.join(os.path.dirname(leon_file), theirs[0]), "rb").read()
    return st


@pytest.mark.parametrize("name", ["fa_k31_blocks", "fq_k31_lossy", "fq_k31_lossless", "fa_k63", "fa_k31_edge"])
def test_headerless_containers_from_our_writer(tmp_path, name):
    import leon_b200
    c = load_case(name)
    if ref_leon(c["k"]) is None:
        pytest.skip("oracle/_ref not built")
    d = tmp_path / "w"
    d.mkdir()
    lf = str(d / ("in.%s.leon" % ("fastq" if c["has_qual"] else "fasta")))
    leon_b200.write_container(lf, result_from_golden(c), c["k"], not c["has_qual"])
    st = _ours_vs_reference(tmp_path, lf, c["k"])
    assert st["nreads"] == len(c["offsets"]) - 1 and not st["has_header"]


@pytest.mark.parametrize("name", ["illumina_fa", "sra_fq", "zeros_fa", "long_fa"])
def test_containers_with_headers_from_our_writer(tmp_path, name):
    import leon_b200
    c = load_hdr(name)
    if ref_leon(c["k"]) is None:
        pytest.skip("oracle/_ref not built")
    d = tmp_path / "w"
    d.mkdir()
    h = leon_b200.encode_headers(c["headers"], c["header_offsets"], c["rpb"], 0)
    lf = str(d / ("in.%s.leon" % ("fastq" if c["has_qual"] else "fasta")))
    leon_b200.write_container(lf, result_from_golden(c), c["k"], not c["has_qual"], header=h)
    st = _ours_vs_reference(tmp_path, lf, c["k"])
    assert st["has_header"] and st["nreads"] == len(c["offsets"]) - 1


@pytest.mark.parametrize("fastq", [True, False])
def test_reference_written_container_through_the_cli(tmp_path, fastq):
    """The reference compresses (chunked HDF5, with names), OUR command line tool decompresses: same bytes as the
    reference's own decoder, and the reads, names and (lossless) qualities of the input."""
    if ref_leon(31) is None:
        pytest.skip("oracle/_ref not built")
    from leon_b200 import synth
    g = synth.make_genome(30000, 8)
    n = 9000
    bases, offs, quals = synth.sample_reads(g, n, 100, 9, n_rate=0.0005, with_quals=fastq)
    d = tmp_path / "r"
    d.mkdir()
    ext = "fastq" if fastq else "fasta"
    inp = d / ("in." + ext)
    o = offs.astype(np.int64)
    with open(inp, "wb") as f:
        for i in range(n):
            s = bases[o[i]:o[i + 1]].tobytes()
            if fastq:
                f.write(b"@SRR.%d lane%d/1\n%s\n+\n%s\n" % (i, i % 3, s, quals[o[i]:o[i + 1]].tobytes()))
            else:
                f.write(b">SRR.%d lane%d/1\n%s\n" % (i, i % 3, s))
    res = subprocess.run([ref_leon(31), "-file", str(inp), "-c", "-nb-cores", "1", "-reads", "2000", "-lossless"], cwd=str(d),
                         capture_output=True, text=True, timeout=900)
    assert res.returncode == 0
    lf = str(inp) + ".leon"
    out = tmp_path / "cli.out"
    res = subprocess.run([CLI, "-file", lf, "-d", "-out", str(out), "-nb-cores", "3"], capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout + res.stderr
    assert open(out, "rb").read() == open(inp, "rb").read()       # lossless, single-line records: the input itself
    res = subprocess.run([ref_leon(31), "-file", lf, "-d", "-nb-cores", "1"], cwd=str(d), capture_output=True, text=True, timeout=900)
    assert res.returncode == 0
    assert open(out, "rb").read() == open(d / ("in.%s.d" % ext), "rb").read()
