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
