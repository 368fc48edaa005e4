"""`.leon` container writer (include/leon_b200.h leon_write_container) on CPU: the datasets the
reference stored for the golden cases (and oracle outputs with many blocks) are written with OUR writer and must
(1) parse as well-formed HDF5 structures holding exactly those bytes and attributes, and (2) be decoded by the compiled
reference (`oracle/_ref/leon -d`) back to the input reads — north_star's third correctness check, minus the GPU."""
import os
import zlib

import numpy as np
import pytest

from container_util import H5Mini, reads_of, ref_leon, reference_decode, result_from_golden
from golden_util import case_names, load_case


def _write(tmp_path, result, k, fastq, name="in"):
    import leon_b200
    lf = str(tmp_path / ("%s.%s.leon" % (name, "fastq" if fastq else "fasta")))
    leon_b200.write_container(lf, result, k, not fastq)
    return lf


def _check_structure(lf, r, k, fastq):
    h = H5Mini(lf).root
    assert sorted(x for x in h if not x.startswith("_")) == ["leon", "metadata"]
    md, leon = h["metadata"], h["leon"]
    assert md["infobyte"]["data"].tolist() == [2 if fastq else 3]
    assert md["infobyte"]["attrs"] == {"type": "fastq" if fastq else "fasta", "header": "false", "version": "1.1.0"}
    assert md["kmerSize"]["data"].view(np.uint64)[0] == k
    assert md["readcount"]["data"].view(np.uint64)[0] == r["read_count"]
    assert md["totalDnaSize"]["data"].view(np.uint64)[0] == r["total_dna_size"]
    assert md["minSequenceSize"]["data"].view(np.int32)[0] == r["min_read_size"]
    assert md["maxSequenceSize"]["data"].view(np.int32)[0] == r["max_read_size"]
    assert np.array_equal(leon["bloom"]["data"], r["bloom"])
    assert leon["bloom"]["attrs"] == {"size": str(int(r["bloom_bits"])), "nb_hash": "7", "type": "neighbor", "kmer_size": str(k)}
    an = leon["anchors"]
    assert np.array_equal(an["anchorsDict"]["data"], r["anchors_dict"])
    assert an["size"]["data"].view(np.uint64)[0] == len(r["anchors_dict"])
    assert an["anchorAdress"]["data"].view(np.uint32)[0] == r["anchor_count"]
    bs = np.asarray(r["blocksizes"], np.uint64)
    nblocks = len(bs) // 2
    dna = leon["dna"]
    assert dna["nb_blocks"]["data"].view(np.uint64)[0] == 2 * nblocks
    assert np.array_equal(dna["blocksizes"]["data"].view(np.uint64), bs)
    at = 0
    for b in range(nblocks):
        n = int(bs[2 * b])
        d = dna["dna_%d" % b]
        assert np.array_equal(d["data"], r["dna"][at:at + n]) and d["attrs"] == {"size": str(n)}
        at += n
    assert [x for x in leon["header"] if not x.startswith("_")] == []
    if fastq:
        qat = 0
        for b in range(nblocks):
            n = int(r["qual_block_sizes"][b])
            q = leon["qual"]["qual_%d" % b]
            assert q["attrs"] == {"size": str(len(q["data"]))}
            assert zlib.decompress(q["data"].tobytes()) == r["qual_text"][qat:qat + n].tobytes()
            qat += n
    else:
        assert "qual" not in leon


@pytest.mark.parametrize("name", case_names())
def test_golden_datasets_round_trip_through_container(tmp_path, name):
    c = load_case(name)
    r = result_from_golden(c)
    lf = _write(tmp_path, r, c["k"], c["has_qual"])
    _check_structure(lf, r, c["k"], c["has_qual"])
    if ref_leon(c["k"]) is None:
        pytest.skip("oracle/_ref not built (needs /root/reference): structure checked, reference decode skipped")
    reads, quals = reference_decode(lf, c["k"], c["has_qual"])
    assert reads == reads_of(c["bases"], c["offsets"])
    if c["has_qual"] and c["lossless"]:
        assert quals == reads_of(c["quals"], c["offsets"])


@pytest.mark.parametrize("nreads,rpb,fastq", [(3000, 10, False), (3000, 1, True), (7, 50000, False)])
def test_many_blocks_multi_level_group_trees(tmp_path, nreads, rpb, fastq):
    """300 and 3000 blocks: the dna/qual groups need 2- and 3-level B-trees of symbol nodes."""
    from leon_b200 import synth
    from oracle import oracle_py as O
    O.build()
    g = synth.make_genome(20000, 3)
    bases, offs, quals = synth.sample_reads(g, nreads, 80, 5, n_rate=0.001, with_quals=fastq)
    r = O.compress(bases, offs, quals, kmer_size=31, reads_per_block=rpb)
    lf = _write(tmp_path, r, 31, fastq)
    _check_structure(lf, r, 31, fastq)
    if ref_leon(31) is None:
        pytest.skip("oracle/_ref not built: structure checked, reference decode skipped")
    reads, _ = reference_decode(lf, 31, fastq)
    assert reads == reads_of(bases, offs)


def test_container_errors(tmp_path):
    import leon_b200
    from leon_b200.api import LeonError
    c = load_case("toy_k31")
    r = result_from_golden(c)
    with pytest.raises(LeonError) as e:
        leon_b200.write_container(str(tmp_path / "no_such_dir" / "x.leon"), r, 31, True)
    assert e.value.code == -6
    with pytest.raises(LeonError) as e:
        leon_b200.write_container(str(tmp_path / "x.leon"), r, 99, True)
    assert e.value.code == -1
    bad = dict(r, blocksizes=np.array([10 ** 9, 5], np.uint64))
    with pytest.raises(LeonError) as e:
        leon_b200.write_container(str(tmp_path / "y.leon"), bad, 31, True)
    assert e.value.code == -1
    assert not os.path.exists(str(tmp_path / "y.leon"))
