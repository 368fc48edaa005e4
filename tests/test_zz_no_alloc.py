"""The entry point that allocates nothing (leon_kmer_offsets_noalloc, include/leon_b200.h; no counterpart in the reference):
called through the C ABI, twice in a row, its output must equal what the
existing entry points (leon_upload + leon_count_resident, read back with leon_debug_fetch "koff") compute -- bit-exact, the
bar of every integer path here.  In a file of its own that pytest collects last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _read_set(seed, nreads, k):
    """Variable-length reads, some shorter than k (they have no k-mer slot)."""
    from leon_b200 import synth
    rng = np.random.Generator(np.random.PCG64(seed))
    lens = rng.integers(1, 180, nreads).astype(np.int64)
    short = rng.random(nreads) < 0.05
    lens[short] = rng.integers(1, k, int(short.sum()))
    offs = np.zeros(nreads + 1, np.uint64)
    np.cumsum(lens, out=offs[1:])
    bases = synth.ACGT[rng.integers(0, 4, int(offs[-1]), dtype=np.uint8)]
    return bases, offs


def _existing(bases, offs, k):
    import leon_b200
    eng = leon_b200.Leon(kmer_size=k, reads_per_block=1000)
    eng.upload(bases, offs)
    eng.count()
    koff = eng.debug_fetch("koff", np.uint64)
    eng.close()
    return koff


def _noalloc(L, offs, k, stream=None):
    import torch
    d_offs = torch.from_numpy(offs.astype(np.int64)).cuda()          # the CALLER owns inputs and outputs (here: torch)
    d_koff = torch.full((len(offs),), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before = torch.cuda.mem_get_info()[0]
    rc = L.leon_kmer_offsets_noalloc(0, d_offs.data_ptr(), len(offs) - 1, k, d_koff.data_ptr(), stream)
    assert rc == 0, L.leon_last_error(None).decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert torch.cuda.mem_get_info()[0] >= free_before - (1 << 20)   # nothing was taken from the device by the call
    return d_koff.cpu().numpy().astype(np.uint64)


@pytest.mark.parametrize("k", [31, 63])
def test_no_alloc_entry_point_equals_existing_path_twice_in_a_row(k):
    import leon_b200
    L = leon_b200.load_library()
    # two calls in a row on the shared scratch: a large set first, then a smaller one (stale tile sums must not leak in),
    # then the large one again
    sets = [_read_set(101, 300_000, k), _read_set(102, 7_001, k)]
    want = [_existing(b, o, k) for b, o in sets]
    for i in (0, 1, 0, 1):
        got = _noalloc(L, sets[i][1], k)
        assert got.shape == want[i].shape
        assert np.array_equal(got, want[i]), "call on set %d differs" % i
        lens = np.diff(sets[i][1].astype(np.int64))
        assert int(got[-1]) == int(np.maximum(lens - k + 1, 0).sum())


def test_no_alloc_entry_point_limits_and_errors():
    import torch
    import leon_b200
    L = leon_b200.load_library()
    d = torch.zeros(8, dtype=torch.int64, device="cuda")
    too_many = (1 << 24) + 1
    rc = L.leon_kmer_offsets_noalloc(0, d.data_ptr(), too_many, 31, d.data_ptr(), None)
    assert rc == -5                                                   # LEON_E_OVERFLOW
    msg = L.leon_last_error(None).decode()
    assert "16777216" in msg and "exceed" in msg
    assert L.leon_kmer_offsets_noalloc(0, None, 4, 31, d.data_ptr(), None) == -1
    assert L.leon_kmer_offsets_noalloc(0, d.data_ptr(), 4, 64, d.data_ptr(), None) == -1
    # an empty set is fine: koff = [0]
    out = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    assert L.leon_kmer_offsets_noalloc(0, d.data_ptr(), 0, 31, out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert int(out.cpu()[0]) == 0
