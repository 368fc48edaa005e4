"""Pins the oracle (oracle/leon_oracle.c) against the compiled reference: golden fixtures produced by
`leon -c -nb-cores 1` (tests/golden/make_golden.py) and the known-answer vectors of SURVEY.md section 8c."""
import ctypes as C

import numpy as np
import pytest

from golden_util import case_names, load_case
from oracle import oracle_py as O


def test_kat_hashes_k31():
    L = O.lib()
    K = 0x0123456789ABCDEF & ((1 << 62) - 1)
    assert L.lo_bloom_seed(0) == 0xffaa54ffe6e6e6e7
    assert L.lo_hash1(O.u128(K), 31, 0) == 0x2e1f57edd3a71e9c
    exp = [0x1572c0763eb01dd8, 0x4a29cb2955bf8cb3, 0x01365b4875557734, 0x29247571345f7430, 0x2cb7d10666ddf5c3,
           0x2395077e09bb2236]
    assert [L.lo_simplehash16(O.u128(K), 31, i) for i in range(1, 7)] == exp
    assert O.from_u128(L.lo_revcomp(O.u128(K), 31)) == 0x147650321cfed8ba
    s = b"GCTCTAGATTTTACAGTTTCTAGTTAGGGGC"
    code = O.from_u128(L.lo_encode_kmer(s, 31))
    assert code == 0x36632a84ea63a3fd
    assert O.from_u128(L.lo_revcomp(O.u128(code), 31)) == 0x355818c06e0098cd


def test_kat_hashes_k63():
    L = O.lib()
    s = b"GCTCTAGATTTTACAGTTTCTAGTTAGGGGCAGCTCATATCTTTACTGATCGCCAGGGAGAGG"
    x = O.from_u128(L.lo_encode_kmer(s, 63))
    assert x == (0x36632a84ea63a3fd << 64) | 0x36489a86c9d4fccf
    assert O.from_u128(L.lo_revcomp(O.u128(x), 63)) == (0x16656f724e0322cd << 64) | 0xb55818c06e0098cd
    assert L.lo_hash1(O.u128(x), 63, 0) == 0x1a312492fdc3cb28
    exp = [0x7d1a76846e051e0e, 0x6e92fd313e581873, 0x2bf6d5a1439c275e, 0x04889c61522c6d4e, 0x1139695b00da6461,
           0x2bf371cb6e430159]
    assert [L.lo_simplehash16(O.u128(x), 63, i) for i in range(1, 7)] == exp


def _bits(b):
    arr = np.ctypeslib.as_array(b.bits, shape=(b.nbytes,))
    return set(np.nonzero(np.unpackbits(arr, bitorder="little"))[0].tolist())


def test_kat_bloom_k31():
    L = O.lib()
    b = O.Bloom()
    assert L.lo_bloom_init(C.byref(b), 2004, 31, 7) == 0
    assert b.nbytes == 1275 and b.reduced_tai == 2004
    s = b"GCTCTAGATTTTACAGTTTCTAGTTAGGGGC"
    x = O.from_u128(L.lo_encode_kmer(s, 31))
    rc = O.from_u128(L.lo_revcomp(O.u128(x), 31))
    L.lo_bloom_insert(C.byref(b), O.u128(min(x, rc)))
    assert _bits(b) == {1820, 2416, 2640, 2857, 3259, 4392, 5432}
    assert L.lo_bloom_contains(C.byref(b), O.u128(min(x, rc)))
    # contains4 re-creates the inserted k-mer: right extension of AGCTC...GGGG by C, left extension of CTCT...GCA by G
    r = L.lo_bloom_contains4(C.byref(b), L.lo_encode_kmer(b"AGCTCTAGATTTTACAGTTTCTAGTTAGGGG", 31), 1)
    assert r == 1 << 1
    l = L.lo_bloom_contains4(C.byref(b), L.lo_encode_kmer(b"CTCTAGATTTTACAGTTTCTAGTTAGGGGCA", 31), 0)
    assert l == 1 << 3
    L.lo_bloom_free(C.byref(b))


def test_kat_bloom_k63():
    L = O.lib()
    b = O.Bloom()
    assert L.lo_bloom_init(C.byref(b), 1596, 63, 7) == 0
    assert b.nbytes == 1224
    s = b"GCTCTAGATTTTACAGTTTCTAGTTAGGGGCAGCTCATATCTTTACTGATCGCCAGGGAGAGG"
    x = O.from_u128(L.lo_encode_kmer(s, 63))
    rc = O.from_u128(L.lo_revcomp(O.u128(x), 63))
    L.lo_bloom_insert(C.byref(b), O.u128(min(x, rc)))
    assert _bits(b) == {1330, 1675, 2451, 3216, 3839, 3999, 4736}
    L.lo_bloom_free(C.byref(b))


def test_kat_rangecoder():
    L = O.lib()
    buf = (C.c_uint8 * 64)()
    n = L.lo_rangecoder_kat(buf, 64)
    assert bytes(buf[:n]).hex() == "16dcd51bc15c990793871000eb56600000"


@pytest.mark.parametrize("name", case_names())
def test_oracle_matches_reference(name):
    c = load_case(name)
    r = O.compress(c["bases"], c["offsets"], c["quals"] if c["has_qual"] else None, kmer_size=c["k"],
                   reads_per_block=c["rpb"], lossless=c["lossless"], abundance_set=c["abundance_set"])
    if c["auto_cutoff"]:
        assert r["cutoff"] == int(c["cutoff"])
        assert r["nb_solids_cutoff"] == int(c["nb_solid"])
    else:
        assert r["nb_solid_total"] == int(c["nb_solid"])
    assert r["bloom_bits"] == int(c["bloom_bits"])
    assert np.array_equal(r["bloom"], c["bloom"]), "Bloom bit array differs from the reference"
    assert r["anchor_count"] == int(c["anchor_count"])
    assert np.array_equal(r["anchors_dict"], c["anchors_dict"])
    assert np.array_equal(r["blocksizes"], c["blocksizes"])
    assert np.array_equal(r["dna"], c["dna_concat"])
    if c["has_qual"]:
        assert np.array_equal(r["qual_block_sizes"], c["qual_text_sizes"])
        assert np.array_equal(r["qual_text"], c["qual_text_concat"])
    assert r["read_count"] == int(c["meta_readcount"])
    assert r["total_dna_size"] == int(c["meta_totalDnaSize"])
    assert r["min_read_size"] == int(c["meta_minSequenceSize"])
    assert r["max_read_size"] == int(c["meta_maxSequenceSize"])
