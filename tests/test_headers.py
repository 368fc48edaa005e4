"""Host header coder (include/leon_b200.h leon_headers_encode) against the reference: the header
streams the compiled reference wrote with -nb-cores 1 (tests/golden/hdrcase_*.npz, made by make_golden_headers.py) must
be reproduced byte for byte whatever the thread count, and a complete container (reference sequence datasets + OUR
header stream, written by OUR container writer) must decode in the reference to the input names and reads."""
import glob
import os
import subprocess

import numpy as np
import pytest

from container_util import reads_of, ref_leon, result_from_golden
from golden_util import GOLDEN


def hdr_cases():
    return sorted(os.path.basename(p)[8:-4] for p in glob.glob(os.path.join(GOLDEN, "hdrcase_*.npz")))


def load_hdr(name):
    z = np.load(os.path.join(GOLDEN, "hdrcase_%s.npz" % name))
    c = {k: z[k] for k in z.files}
    c["k"], c["rpb"], c["has_qual"] = int(c["kmer_size"]), int(c["reads_per_block"]), len(c["quals"]) > 0
    return c


@pytest.mark.parametrize("nthreads", [1, 3, 0])
@pytest.mark.parametrize("name", hdr_cases())
def test_header_stream_equals_reference(name, nthreads):
    import leon_b200
    c = load_hdr(name)
    h = leon_b200.encode_headers(c["headers"], c["header_offsets"], c["rpb"], nthreads)
    assert h["first_header"] == c["first_header"].tobytes()
    assert np.array_equal(h["blocksizes"], c["header_blocksizes"])
    assert np.array_equal(h["data"], c["header_concat"])


def test_ragged_headers_fall_back_to_the_serial_coder():
    """Numeric fields past the previous header's field count read state the reference never resets: the parallel
    coder must notice and recode serially (and still equal the reference, checked above)."""
    import leon_b200
    c = load_hdr("ragged_fa")
    assert leon_b200.encode_headers(c["headers"], c["header_offsets"], c["rpb"], 4)["coded_serially"]
    c = load_hdr("illumina_fa")
    assert not leon_b200.encode_headers(c["headers"], c["header_offsets"], c["rpb"], 4)["coded_serially"]


def test_header_coder_edge_cases():
    import leon_b200
    from leon_b200.api import LeonError
    h = leon_b200.encode_headers(np.zeros(0, np.uint8), np.zeros(1, np.uint64), 100)
    assert h["first_header"] == b"" and len(h["blocksizes"]) == 0 and len(h["data"]) == 0
    names = [b"a", b"", b"7", b"007", b"a" * 300, b"18446744073709551616", b"x 1", b"x 1"]
    text = np.frombuffer(b"".join(names), np.uint8)
    offs = np.concatenate([[0], np.cumsum([len(n) for n in names])]).astype(np.uint64)
    a = leon_b200.encode_headers(text, offs, 3, 1)
    b = leon_b200.encode_headers(text, offs, 3, 2)
    assert a["blocksizes"].tolist()[1::2] == [3, 3, 2] and np.array_equal(a["data"], b["data"])
    bad = np.frombuffer("réad".encode("latin-1"), np.uint8)
    with pytest.raises(LeonError):
        leon_b200.encode_headers(bad, np.array([0, len(bad)], np.uint64), 10)


@pytest.mark.parametrize("name", hdr_cases())
def test_reference_decodes_container_with_our_header_stream(tmp_path, name):
    import leon_b200
    c = load_hdr(name)
    if not bool(c["ref_roundtrip_names_ok"]):
        pytest.skip("the reference cannot decode its own file for this case (its decoder throws on these headers)")
    if ref_leon(c["k"]) is None:
        pytest.skip("oracle/_ref not built")
    h = leon_b200.encode_headers(c["headers"], c["header_offsets"], c["rpb"], 0)
    ext = "fastq" if c["has_qual"] else "fasta"
    lf = str(tmp_path / ("in.%s.leon" % ext))
    leon_b200.write_container(lf, result_from_golden(c), c["k"], not c["has_qual"], header=h)
    res = subprocess.run([ref_leon(c["k"]), "-file", lf, "-d", "-nb-cores", "1"], cwd=str(tmp_path), capture_output=True,
                         text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-1500:] + res.stderr[-1500:]
    lines = [l.rstrip(b"\n") for l in open(str(tmp_path / ("in.%s.d" % ext)), "rb")]
    step = 4 if c["has_qual"] else 2
    assert [l[1:] for l in lines[0::step]] == reads_of(c["headers"], c["header_offsets"])
    assert lines[1::step] == reads_of(c["bases"], c["offsets"])
