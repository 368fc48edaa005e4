"""The GPU decompressor (leon_decode_dna_device / k_decode_blocks) on the GPU box.
(1) The sequence streams the REFERENCE wrote (golden fixtures) decode on the device to the input reads.
(2) Streams written by our compressor (N runs, indels, variable lengths, k = 31 and k = 63, more blocks than resident
    warps) decode to the input, and to what the host-thread decoder returns.
(3) File level: `leon_b200 -d -gpu-decode` writes byte for byte the file `leon_b200 -d` (host threads) writes.
(4) Corrupt streams end as errors or as some reads, never as a hang or a fault."""
import os
import subprocess

import numpy as np
import pytest

from container_util import reads_of, result_from_golden, ROOT
from golden_util import case_names, load_case

pytestmark = pytest.mark.gpu
CLI = os.path.join(ROOT, "leon_b200", "bin", "leon_b200")


@pytest.mark.parametrize("name", case_names())
def test_reference_sequence_streams_decode_on_the_device(name):
    import leon_b200
    c = load_case(name)
    r = result_from_golden(c)
    assert leon_b200.decode_dna_device(r, c["k"]) == reads_of(c["bases"], c["offsets"])


@pytest.mark.parametrize("k,read_len,nreads,rpb,indels", [(31, 100, 150000, 100, False), (31, 150, 60000, 2000, True),
                                                          (63, 150, 40000, 1500, True), (21, 76, 30000, 50000, False)])
def test_our_streams_round_trip_through_the_device_decoder(k, read_len, nreads, rpb, indels):
    import leon_b200
    from leon_b200 import synth
    g = synth.make_genome(40000, 11)
    bases, offs, _ = synth.sample_reads(g, nreads, read_len, 12, n_rate=0.001)
    if indels:
        bases, offs, _ = synth.add_indels_and_n_runs(bases, offs, None, 13)
    eng = leon_b200.Leon(kmer_size=k, reads_per_block=rpb)
    r = eng.compress(bases, offs)
    eng.close()
    want = reads_of(bases, offs)
    got, ms = leon_b200.decode_dna_device(r, k, with_time=True)
    assert got == want
    assert ms > 0
    assert leon_b200.decode_dna(r, k, 4) == want


def test_file_level_device_decode_equals_host_decode(tmp_path):
    from test_gpu_file import _write_file
    path = str(tmp_path / "in.fastq")
    _write_file(path, True, 23000, 5)
    res = subprocess.run([CLI, "-file", path, "-c", "-kmer-size", "31", "-reads", "5000"], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout + res.stderr
    a, b = str(tmp_path / "host.out"), str(tmp_path / "gpu.out")
    res = subprocess.run([CLI, "-file", path + ".leon", "-d", "-out", a], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout + res.stderr
    res = subprocess.run([CLI, "-file", path + ".leon", "-d", "-gpu-decode", "-out", b], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "k_decode_blocks" in res.stdout
    assert open(a, "rb").read() == open(b, "rb").read()


def test_device_decoder_without_max_read_size_is_an_error():
    import leon_b200
    r = dict(result_from_golden(load_case("toy_k31")))
    r["max_read_size"] = -1
    with pytest.raises(leon_b200.LeonError):
        leon_b200.decode_dna_device(r, 31)


@pytest.mark.timeout(300)
def test_corrupt_streams_on_the_device_are_errors_not_crashes():
    import leon_b200
    c = load_case("fa_k31_blocks")
    r = result_from_golden(c)
    rng = np.random.default_rng(5)
    for trial in range(6):
        bad = dict(r)
        dna = r["dna"].copy()
        pos = rng.integers(0, len(dna), 40)
        dna[pos] = rng.integers(0, 256, 40)
        bad["dna"] = dna
        try:
            leon_b200.decode_dna_device(bad, c["k"])
        except leon_b200.LeonError:
            pass
    # the device is still usable afterwards
    assert leon_b200.decode_dna_device(r, c["k"]) == reads_of(c["bases"], c["offsets"])
