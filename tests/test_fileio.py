"""FASTA/FASTQ record reader of the file-level path (leon_read_file) on CPU: layouts the reference's bank accepts
(bank/impl/BankFasta.cpp:484-570): multi-line sequences, blank lines, CR LF, text before the first record, '@' inside
qualities, multi-line FASTQ, a missing final newline."""
import numpy as np
import pytest


def _read(tmp_path, name, text, **kw):
    import leon_b200
    p = tmp_path / name
    p.write_bytes(text)
    return leon_b200.read_file(str(p), **kw)


def _split(a, offs):
    o = offs.astype(np.int64)
    return [a[o[i]:o[i + 1]].tobytes() for i in range(len(o) - 1)]


def test_fasta_layouts(tmp_path):
    r = _read(tmp_path, "a.fa", b"junk line\n>r1 first read\nACGT\nNNAC\n\n>r2\r\nGG\r\nTT\r\n>\n>r4\nA", want_quals=False)
    assert _split(r["bases"], r["offsets"]) == [b"ACGTNNAC", b"GGTT", b"", b"A"]
    assert _split(r["headers"], r["header_offsets"]) == [b"r1 first read", b"r2", b"", b"r4"]
    assert not r["is_fastq"] and r["quals"] is None


def test_fastq_layouts(tmp_path):
    txt = b"@r1 x\nACGT\n+\n@II>\n@r2\nAC\nGT\n+r2\n!!\n+@\n@r3\n\n+\n\n@r4\nT\n+\n#"
    r = _read(tmp_path, "a.fq", txt)
    assert _split(r["bases"], r["offsets"]) == [b"ACGT", b"ACGT", b"", b"T"]
    assert _split(r["quals"], r["offsets"]) == [b"@II>", b"!!+@", b"", b"#"]
    assert _split(r["headers"], r["header_offsets"]) == [b"r1 x", b"r2", b"r3", b"r4"]
    assert r["is_fastq"]
    r = _read(tmp_path, "a.fq", txt, want_quals=False, want_headers=False)
    assert _split(r["bases"], r["offsets"]) == [b"ACGT", b"ACGT", b"", b"T"] and r["headers"] is None


def test_reader_errors(tmp_path):
    import leon_b200
    from leon_b200.api import LeonError
    with pytest.raises(LeonError) as e:
        leon_b200.read_file(str(tmp_path / "missing.fa"))
    assert e.value.code == -6
    with pytest.raises(LeonError):
        _read(tmp_path, "b.fq", b"@r1\nACGT\n+\nIIIIII\n")          # quality longer than the sequence
    with pytest.raises(LeonError):
        _read(tmp_path, "c.fq", b"@r1\nACGT\n+\nII\n")              # quality shorter than the sequence
    with pytest.raises(LeonError):
        _read(tmp_path, "d.fq", b">r1\nACGT\n")                      # qualities wanted, FASTA given
    r = _read(tmp_path, "e.fa", b"")
    assert len(r["offsets"]) == 1
    with pytest.raises(LeonError) as e:
        leon_b200.compress_file(str(tmp_path / "reads.txt"))
    assert e.value.code == -1 and "extension" in str(e.value)


def _big_records(n, seed, fastq):
    rng = np.random.Generator(np.random.PCG64(seed))
    acgt = np.frombuffer(b"ACGTN", np.uint8)
    names, reads, quals = [], [], []
    for i in range(n):
        L = int(rng.integers(60, 140))
        reads.append(acgt[rng.integers(0, 5, L)].tobytes())
        names.append(b"read.%d lane:%d" % (i, i % 7))
        # qualities over the whole printable range: lines that begin with '@', '+' or '>' are common
        quals.append(rng.integers(33, 75, L).astype(np.uint8).tobytes() if fastq else None)
    return names, reads, quals


@pytest.mark.parametrize("fastq", [True, False])
def test_parallel_reader_equals_serial(tmp_path, fastq):
    """Files above the 4 MB threshold are cut at record starts and read by several threads."""
    import leon_b200
    names, reads, quals = _big_records(45000, 3, fastq)
    assert sum(q[:1] in (b"@", b"+", b">") for q in quals if q) > 1000 or not fastq
    p = tmp_path / ("big.fq" if fastq else "big.fa")
    with open(p, "wb") as f:
        for i in range(len(reads)):
            if fastq:
                f.write(b"@" + names[i] + b"\n" + reads[i] + b"\n+\n" + quals[i] + b"\n")
            else:
                r = reads[i]
                f.write(b">" + names[i] + b"\n" + b"\n".join(r[j:j + 50] for j in range(0, len(r), 50)) + b"\n")
    r = leon_b200.read_file(str(p), want_quals=fastq)
    assert r["threads_used"] > 1
    assert _split(r["bases"], r["offsets"]) == reads
    assert _split(r["headers"], r["header_offsets"]) == names
    if fastq:
        assert _split(r["quals"], r["offsets"]) == quals


def test_irregular_big_fastq_falls_back_to_one_thread(tmp_path):
    """Multi-line FASTQ cannot be cut safely: the reader must notice and parse it serially, with the same result."""
    import leon_b200
    names, reads, quals = _big_records(45000, 4, True)
    p = tmp_path / "multi.fastq"
    with open(p, "wb") as f:
        for i in range(len(reads)):
            r, q = reads[i], quals[i]
            if i % 1000 == 999:      # now and then a record with 2-line sequence and quality
                h = len(r) // 2
                f.write(b"@" + names[i] + b"\n" + r[:h] + b"\n" + r[h:] + b"\n+\n" + q[:h] + b"\n" + q[h:] + b"\n")
            else:
                f.write(b"@" + names[i] + b"\n" + r + b"\n+\n" + q + b"\n")
    r = leon_b200.read_file(str(p))
    assert r["threads_used"] == 1
    assert _split(r["bases"], r["offsets"]) == reads and _split(r["quals"], r["offsets"]) == quals
    assert _split(r["headers"], r["header_offsets"]) == names


def test_gzip_input_is_inflated(tmp_path):
    """The reference reads its input through gzopen (bank/impl/BankFasta.cpp:652): a .gz file parses like its text, also
    when it is a concatenation of gzip members (bgzip), and a truncated one is an error, not a bogus read set."""
    import gzip
    import leon_b200
    from leon_b200.api import LeonError
    txt = b"".join(b"@r%d\nACGTNACGTTGCA%s\n+\nIIIIIIIIIIIII%s\n" % (i, b"ACGT"[i % 4:], b"#!+@"[i % 4:]) for i in range(3000))
    plain = _read(tmp_path, "a.fastq", txt)
    (tmp_path / "b.fastq.gz").write_bytes(gzip.compress(txt))
    gz = leon_b200.read_file(str(tmp_path / "b.fastq.gz"))
    half = len(txt) // 2
    cut = txt.rfind(b"\n@r", 0, half) + 1
    (tmp_path / "c.fastq.gz").write_bytes(gzip.compress(txt[:cut]) + gzip.compress(txt[cut:]))
    gz2 = leon_b200.read_file(str(tmp_path / "c.fastq.gz"))
    for r in (gz, gz2):
        for key in ("bases", "offsets", "quals", "headers", "header_offsets"):
            assert np.array_equal(r[key], plain[key]), key
    (tmp_path / "d.fastq.gz").write_bytes(gzip.compress(txt)[:-40])
    with pytest.raises(LeonError):
        leon_b200.read_file(str(tmp_path / "d.fastq.gz"))


def test_workload_writer_and_digest(tmp_path):
    """leon_b200.workloads: the vectorised file writer equals the per-read one, chunk-seeded generation is independent of
    which chunks are asked for, and the digest tells streams apart."""
    from leon_b200 import synth, workloads as W
    g = synth.make_genome(5000, 1)
    b, o, q = synth.sample_reads(g, 1234, 100, 2, with_quals=True)
    W.write_fixed(str(tmp_path / "a.fq"), b, 100, q)
    synth.write_fasta(str(tmp_path / "b.fq"), b, o, q)
    assert (tmp_path / "a.fq").read_bytes() == (tmp_path / "b.fq").read_bytes()
    W.write_fixed(str(tmp_path / "a.fa"), b, 100)
    synth.write_fasta(str(tmp_path / "b.fa"), b, o)
    assert (tmp_path / "a.fa").read_bytes() == (tmp_path / "b.fa").read_bytes()
    n = W.C3_CHUNK + 1000
    fb, fo, fq = W.generate_c3(300000, n)
    cb, co, cq = W.generate_c3(300000, n, chunks=[1])
    assert len(co) - 1 == 1000 and np.array_equal(cb, fb[int(fo[W.C3_CHUNK]):]) and np.array_equal(cq, fq[int(fo[W.C3_CHUNK]):])
    lens = np.diff(fo.astype(np.int64))
    assert lens.max() == 150 and lens.min() < 150 and (fb == ord("N")).any()
    r = dict(bloom=np.arange(10, dtype=np.uint8), anchors_dict=np.zeros(3, np.uint8), dna=np.ones(7, np.uint8),
             blocksizes=np.array([7, 2], np.uint64), qual_text=np.frombuffer(b"II\n#I\n", np.uint8), qual_block_sizes=np.array([6], np.uint64))
    d1 = W.digest_of_result(r, True)
    r2 = dict(r, dna=np.zeros(7, np.uint8))
    ok, diff = W.digests_equal(d1, W.digest_of_result(r2, True))
    assert not ok and diff == ["dna"]
    assert W.digests_equal(d1, W.digest_of_result(dict(r), True))[0]
