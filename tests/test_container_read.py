"""`.leon` container reader (leon_container_*): files written by leon_write_container and, where the compiled reference is
available, files the REFERENCE wrote through the HDF5 library (chunked datasets) must give back the same datasets and
attributes as the reference's own gatb-h5dump."""
import os
import subprocess

import numpy as np
import pytest

from container_util import ref_attr, ref_dataset, ref_h5dump, ref_leon, result_from_golden
from golden_util import load_case


def test_reads_back_what_the_writer_wrote(tmp_path):
    import leon_b200
    c = load_case("fq_k31_lossy")
    r = result_from_golden(c)
    lf = str(tmp_path / "in.fastq.leon")
    leon_b200.write_container(lf, r, 31, False)
    with leon_b200.Container(lf) as h:
        assert h.has("leon/dna") and h.has("leon/qual/qual_0") and not h.has("leon/header/firstheader") and not h.has("nope/x")
        assert np.array_equal(h.dataset("leon/bloom"), r["bloom"])
        assert h.attr("leon/bloom", "size") == str(r["bloom_bits"]) and h.attr("leon/bloom", "type") == "neighbor"
        assert h.attr("metadata/infobyte", "type") == "fastq" and h.attr("metadata/infobyte", "version") == "1.1.0"
        assert np.array_equal(h.dataset("leon/anchors/anchorsDict"), r["anchors_dict"])
        assert np.array_equal(h.dataset("/leon/dna/blocksizes").view(np.uint64), r["blocksizes"])
        at = 0
        for b in range(len(r["blocksizes"]) // 2):
            n = int(r["blocksizes"][2 * b])
            assert np.array_equal(h.dataset("leon/dna/dna_%d" % b), r["dna"][at:at + n])
            at += n
        with pytest.raises(leon_b200.LeonError):
            h.dataset("leon/dna/dna_999")
        with pytest.raises(leon_b200.LeonError):
            h.attr("leon/bloom", "nope")


def test_rejects_other_files(tmp_path):
    import leon_b200
    p = tmp_path / "x.leon"
    p.write_bytes(b"not an hdf5 file" * 10)
    with pytest.raises(leon_b200.LeonError):
        leon_b200.Container(str(p))
    with pytest.raises(leon_b200.LeonError):
        leon_b200.Container(str(tmp_path / "missing.leon"))


@pytest.mark.parametrize("fastq,nreads,rpb", [(True, 3000, 700), (False, 40000, 50000)])
def test_reads_reference_written_containers(tmp_path, fastq, nreads, rpb):
    """Chunked HDF5 as the reference writes it (with read names: header group filled); the larger case makes the
    dna_0 dataset span several chunks of its B-tree."""
    if ref_leon(31) is None or ref_h5dump() is None:
        pytest.skip("oracle/_ref not built")
    import leon_b200
    from leon_b200 import synth
    g = synth.make_genome(30000, 5)
    bases, offs, quals = synth.sample_reads(g, nreads, 100, 6, n_rate=0.0005, with_quals=fastq)
    ext = "fastq" if fastq else "fasta"
    inp = tmp_path / ("in." + ext)
    o = offs.astype(np.int64)
    with open(inp, "wb") as f:
        for i in range(nreads):
            s = bases[o[i]:o[i + 1]].tobytes()
            if fastq:
                f.write(b"@read%d/1\n%s\n+\n%s\n" % (i, s, quals[o[i]:o[i + 1]].tobytes()))
            else:
                f.write(b">read%d/1\n%s\n" % (i, s))
    res = subprocess.run([ref_leon(31), "-file", str(inp), "-c", "-nb-cores", "1", "-reads", str(rpb)], cwd=str(tmp_path),
                         capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-1500:] + res.stderr[-1500:]
    lf = str(inp) + ".leon"
    nblocks = (nreads + rpb - 1) // rpb
    names = ["metadata/infobyte", "metadata/kmerSize", "metadata/readcount", "leon/bloom", "leon/anchors/anchorsDict",
             "leon/anchors/size", "leon/anchors/anchorAdress", "leon/dna/nb_blocks", "leon/dna/blocksizes",
             "leon/header/firstheader", "leon/header/blocksizes"]
    names += ["leon/dna/dna_%d" % b for b in range(nblocks)] + ["leon/header/header_%d" % b for b in range(nblocks)]
    if fastq:
        names += ["leon/qual/qual_%d" % b for b in range(nblocks)]
    with leon_b200.Container(lf) as h:
        for nm in names:
            assert np.array_equal(h.dataset(nm), ref_dataset(lf, "/" + nm, str(tmp_path))), nm
        for a in ("size", "nb_hash", "type", "kmer_size"):
            assert h.attr("leon/bloom", a) == ref_attr(lf, "/leon/bloom/" + a)
        for a in ("type", "header", "version"):
            assert h.attr("metadata/infobyte", a) == ref_attr(lf, "/metadata/infobyte/" + a)
        assert h.attr("leon/dna/dna_0", "size") == ref_attr(lf, "/leon/dna/dna_0/size")
