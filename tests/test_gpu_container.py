"""GPU -> `.leon` container -> compiled reference decoder (north_star's third correctness check): reads compressed by
the CUDA path and written with leon_write_container must come back from `oracle/_ref/leon -d` as the input reads
(and, with -lossless, the input qualities)."""
import numpy as np
import pytest

from container_util import reads_of, ref_leon, reference_decode
from golden_util import load_case

pytestmark = pytest.mark.gpu


def _round_trip(tmp_path, bases, offs, quals, k, rpb, lossless=False):
    import leon_b200
    fastq = quals is not None
    eng = leon_b200.Leon(kmer_size=k, reads_per_block=rpb, has_qual=fastq, lossless=lossless)
    r = eng.compress(bases, offs, quals)
    eng.close()
    lf = str(tmp_path / ("in.%s.leon" % ("fastq" if fastq else "fasta")))
    leon_b200.write_container(lf, r, k, not fastq)
    if ref_leon(k) is None:
        pytest.skip("oracle/_ref not built")
    reads, q = reference_decode(lf, k, fastq)
    assert reads == reads_of(bases, offs)
    if fastq and lossless:
        assert q == reads_of(quals, offs)
    return r


@pytest.mark.parametrize("name", ["fa_k31_blocks", "fa_k31_edge", "fq_k31_lossy", "fq_k31_lossless", "fa_k63", "fq_k47_lossy"])
def test_reference_decodes_gpu_output_of_golden_inputs(tmp_path, name):
    c = load_case(name)
    _round_trip(tmp_path, c["bases"], c["offsets"], c["quals"] if c["has_qual"] else None, c["k"], c["rpb"], c["lossless"])


@pytest.mark.parametrize("fastq", [False, True])
def test_reference_decodes_gpu_output_400k_reads(tmp_path, fastq):
    """8 blocks of 50000 reads with substitutions, indels and N runs (default block size of the reference)."""
    from leon_b200 import synth
    g = synth.make_genome(400000, 11)
    bases, offs, quals = synth.sample_reads(g, 400000, 100, 12, n_rate=0.0005, with_quals=fastq)
    bases, offs, quals = synth.add_indels_and_n_runs(bases, offs, quals, 13, 0.0005, 0.002)
    r = _round_trip(tmp_path, bases, offs, quals, 31, 50000, lossless=fastq)
    assert r["nblocks"] == 8


def _inflate_blocks(r):
    import zlib
    out, at = [], 0
    for n in r["qual_z_block_sizes"]:
        out.append(zlib.decompress(r["qual_z"][at:at + int(n)].tobytes()))
        at += int(n)
    assert at == len(r["qual_z"])
    return out


@pytest.mark.parametrize("name", ["fq_k31_lossy", "fq_k31_lossless", "fq_k47_lossy"])
def test_device_zlib_streams_inflate_to_the_quality_text(name):
    """Quality blocks deflated on the GPU (deflate.cuh): every stream must inflate, with host zlib, to exactly the text
    the reference hands to its own deflate (golden fixture), and must not be much larger than zlib's own output."""
    import zlib
    import leon_b200
    c = load_case(name)
    eng = leon_b200.Leon(kmer_size=c["k"], reads_per_block=c["rpb"], has_qual=True, lossless=c["lossless"], qual_zlib_on_device=True)
    r = eng.compress(c["bases"], c["offsets"], c["quals"])
    eng.close()
    got = _inflate_blocks(r)
    sizes = [int(x) for x in c["qual_text_sizes"]]
    assert [len(g) for g in got] == sizes
    assert b"".join(got) == c["qual_text_concat"].tobytes()
    assert np.array_equal(r["qual_text"], c["qual_text_concat"])          # the text itself is still returned
    ref = sum(len(zlib.compress(g, 9)) for g in got)
    assert len(r["qual_z"]) <= 1.10 * ref + 64 * len(got), (len(r["qual_z"]), ref)


def test_device_zlib_large_blocks_and_skip_text(tmp_path):
    """5 MB blocks (79 chunks each), lossy and lossless; with skip_qual_text only the streams come back, and the
    container written from them decodes in the reference."""
    import zlib
    import leon_b200
    from leon_b200 import synth
    from oracle import oracle_py as O
    g = synth.make_genome(150000, 21)
    bases, offs, quals = synth.sample_reads(g, 120000, 100, 22, n_rate=0.0005, with_quals=True)
    for lossless in (False, True):
        ref = O.compress(bases, offs, quals, kmer_size=31, reads_per_block=50000, lossless=lossless)
        eng = leon_b200.Leon(kmer_size=31, reads_per_block=50000, has_qual=True, lossless=lossless, qual_zlib_on_device=True,
                             skip_qual_text=True)
        r = eng.compress(bases, offs, quals)
        eng.close()
        assert len(r["qual_text"]) == 0 and r["nblocks"] == 3
        assert b"".join(_inflate_blocks(r)) == ref["qual_text"].tobytes()
        z9 = len(zlib.compress(ref["qual_text"].tobytes(), 9))
        assert len(r["qual_z"]) <= 1.10 * z9, (lossless, len(r["qual_z"]), z9)
        if not lossless and ref_leon(31) is not None:
            lf = str(tmp_path / "in.fastq.leon")
            leon_b200.write_container(lf, r, 31, False)
            reads, _ = reference_decode(lf, 31, True)
            assert reads == reads_of(bases, offs)
