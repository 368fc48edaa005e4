"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, must reproduce the
reference byte for byte — against the committed golden fixtures (outputs of the compiled reference) and against the
oracle on seeded inputs — plus size-independent properties at larger sizes."""
import numpy as np
import pytest

from golden_util import case_names, load_case

pytestmark = pytest.mark.gpu


def _engine(c=None, **kw):
    import leon_b200
    if c is not None:
        kw = dict(kmer_size=c["k"], reads_per_block=c["rpb"], lossless=c["lossless"],
                  abundance=3 if c["abundance_set"] else None, has_qual=c["has_qual"], **kw)
    return leon_b200.Leon(**kw)


def _same(r, ref, has_qual):
    for key in ("cutoff", "nb_solids_cutoff", "nb_solid_total", "bloom_bits", "anchor_count", "read_count",
                "total_dna_size", "reads_without_anchor"):
        assert r[key] == ref[key], key
    for key in ("histogram", "bloom", "anchors_dict", "blocksizes", "dna", "anchor_pos"):
        assert np.array_equal(r[key], ref[key]), key
    am = ref["anchor_pos"] >= 0
    assert np.array_equal(r["anchor_addr"][am], ref["anchor_addr"][am])
    if has_qual:
        assert np.array_equal(r["qual_text"], ref["qual_text"])
        assert np.array_equal(r["qual_block_sizes"], ref["qual_block_sizes"])


@pytest.mark.parametrize("name", case_names())
def test_golden_reference_bytes(name):
    """CUDA output == datasets of the .leon file the compiled reference wrote with -nb-cores 1."""
    c = load_case(name)
    eng = _engine(c)
    r = eng.compress(c["bases"], c["offsets"], c["quals"] if c["has_qual"] else None)
    eng.close()
    if c["auto_cutoff"]:
        assert r["cutoff"] == int(c["cutoff"]) and r["nb_solids_cutoff"] == int(c["nb_solid"])
    else:
        assert r["nb_solid_total"] == int(c["nb_solid"])
    assert r["bloom_bits"] == int(c["bloom_bits"])
    assert np.array_equal(r["bloom"], c["bloom"])
    assert r["anchor_count"] == int(c["anchor_count"])
    assert np.array_equal(r["anchors_dict"], c["anchors_dict"])
    assert np.array_equal(r["blocksizes"], c["blocksizes"])
    assert np.array_equal(r["dna"], c["dna_concat"])
    if c["has_qual"]:
        assert np.array_equal(r["qual_block_sizes"], c["qual_text_sizes"])
        assert np.array_equal(r["qual_text"], c["qual_text_concat"])
    assert r["read_count"] == int(c["meta_readcount"]) and r["total_dna_size"] == int(c["meta_totalDnaSize"])
    assert r["min_read_size"] == int(c["meta_minSequenceSize"]) and r["max_read_size"] == int(c["meta_maxSequenceSize"])


@pytest.mark.parametrize("name,batch", [("fa_k31_blocks", 64), ("fa_k31_blocks", 1000), ("fa_k31_edge", 37),
                                        ("fq_k31_lossy", 501), ("fa_k63", 128), ("fa_k31_lowcov", 256)])
def test_anchor_batching_is_invisible(name, batch):
    """Anchor resolution must equal the serial reference whatever the batch size."""
    from oracle import oracle_py as O
    c = load_case(name)
    ref = O.compress(c["bases"], c["offsets"], c["quals"] if c["has_qual"] else None, kmer_size=c["k"],
                     reads_per_block=c["rpb"], lossless=c["lossless"], abundance_set=c["abundance_set"])
    eng = _engine(c, anchor_batch=batch)
    r = eng.compress(c["bases"], c["offsets"], c["quals"] if c["has_qual"] else None)
    eng.close()
    _same(r, ref, c["has_qual"])


@pytest.mark.parametrize("k,L,fastq,seed", [(31, 100, False, 1), (31, 150, True, 2), (25, 76, False, 3), (63, 150, False, 4),
                                            (40, 101, True, 5)])
def test_oracle_parity_seeded(k, L, fastq, seed):
    """Seeded synthetic sets of the BASELINE shape (reduced), several blocks, vs the oracle."""
    from oracle import oracle_py as O
    from leon_b200 import synth
    g = synth.make_genome(60000, seed)
    bases, offs, quals = synth.sample_reads(g, 40000, L, seed + 100, n_rate=0.0005, with_quals=fastq)
    ref = O.compress(bases, offs, quals, kmer_size=k, reads_per_block=7000)
    eng = _engine(kmer_size=k, reads_per_block=7000, has_qual=fastq)
    r = eng.compress(bases, offs, quals)
    eng.close()
    _same(r, ref, fastq)


def test_bloom_queries_match_oracle():
    """contains() and contains4() of the device filter == BloomNeighborCoherent restated, on real and random k-mers."""
    import ctypes as C
    from oracle import oracle_py as O
    for name in ("fa_k31_blocks", "fa_k63"):
        c = load_case(name)
        k = c["k"]
        eng = _engine(c)
        eng.upload(c["bases"], c["offsets"])
        eng.count()
        bloom, bits = eng.bloom_build()
        assert np.array_equal(bloom, c["bloom"])
        L = O.lib()
        b = O.Bloom()
        assert L.lo_bloom_init(C.byref(b), bits, k, 7) == 0   # bits = reduced size: same geometry
        assert b.nbytes == len(bloom)
        C.memmove(b.bits, bloom.ctypes.data, len(bloom))
        rng = np.random.default_rng(7)
        vals = []
        s = c["bases"]
        for _ in range(1500):   # k-mers taken from reads (mostly present) ...
            p = int(rng.integers(0, len(s) - k))
            w = bytes(s[p:p + k]).replace(b"N", b"A")
            vals.append(O.from_u128(L.lo_encode_kmer(w, k)))
        for _ in range(1500):   # ... and random ones (mostly absent)
            vals.append(int.from_bytes(rng.bytes(16), "little") & ((1 << (2 * k)) - 1))
        km = np.array([[v & (2**64 - 1), v >> 64] for v in vals], np.uint64)
        for right in (0, 1):
            o4, o1 = eng.debug_bloom_query(km, right)
            for i, v in enumerate(vals):
                assert int(o4[i]) == L.lo_bloom_contains4(C.byref(b), O.u128(v), right)
                rc = O.from_u128(L.lo_revcomp(O.u128(v), k))
                assert int(o1[i]) == (1 if L.lo_bloom_contains(C.byref(b), O.u128(min(v, rc))) else 0)
        L.lo_bloom_free(C.byref(b))
        eng.close()


@pytest.mark.parametrize("name", ["fa_k31_blocks", "fa_k63", "fq_k31_lossy", "fa_k31_edge"])
def test_dictionary_coder_on_device_and_on_host_agree(name):
    """The anchor dictionary dump is the same bytes whether its serial recurrence runs as a kernel or on the host thread
    (the default), and both equal the reference's anchorsDict."""
    c = load_case(name)
    for on_device in (True, False):
        eng = _engine(c, dict_on_device=on_device)
        r = eng.compress(c["bases"], c["offsets"], c["quals"] if c["has_qual"] else None)
        eng.close()
        assert r["anchor_count"] == int(c["anchor_count"])
        assert np.array_equal(r["anchors_dict"], c["anchors_dict"]), "on_device=%s" % on_device
        assert np.array_equal(r["dna"], c["dna_concat"])


def test_edge_inputs():
    """Empty set, reads shorter than k, all-N reads, a single read: same bytes as the oracle, no faults."""
    from oracle import oracle_py as O
    cases = []
    cases.append((np.zeros(0, np.uint8), np.zeros(1, np.uint64)))
    short = np.frombuffer(b"ACGTACGTAC" * 3, np.uint8)
    cases.append((np.concatenate([short[:10], short[:30], short[:5]]), np.array([0, 10, 40, 45], np.uint64)))
    cases.append((np.frombuffer(b"N" * 80 + b"ACGT" * 20, np.uint8).copy(), np.array([0, 80, 160], np.uint64)))
    one = np.frombuffer(b"GATTACAGATTACAGGATCCGATTACATTTAGGACCAGATTACAGATTACAGGATCCGATTACATTTAGGACCA", np.uint8)
    cases.append((one.copy(), np.array([0, len(one)], np.uint64)))
    for bases, offs in cases:
        ref = O.compress(bases, offs, None, kmer_size=31, reads_per_block=2)
        eng = _engine(kmer_size=31, reads_per_block=2)
        r = eng.compress(bases, offs)
        eng.close()
        _same(r, ref, False)


def test_properties_full_shape_reduced_count():
    """Size-independent properties on a BASELINE-config-2-shaped set too large for byte goldens: block table sums,
    anchors are Bloom-positive canonical k-mers of their read, and the result is reproducible."""
    from leon_b200 import synth
    g = synth.make_genome(1_000_000, 9)
    bases, offs, _ = synth.sample_reads(g, 300_000, 100, 10)
    eng = _engine(kmer_size=31, reads_per_block=50000)
    r1 = eng.compress(bases, offs)
    r2 = eng.compress(bases, offs)
    assert r1["nblocks"] == 6
    assert int(r1["blocksizes"][0::2].sum()) == len(r1["dna"])
    assert int(r1["blocksizes"][1::2].sum()) == 300_000
    for key in ("bloom", "anchors_dict", "dna", "blocksizes", "anchor_pos", "anchor_addr"):
        assert np.array_equal(r1[key], r2[key]), key
    assert r1["anchor_addr"][r1["anchor_pos"] >= 0].max() == r1["anchor_count"] - 1
    # addresses are handed out in read order: the first read using address a comes before the first using a+1
    am = r1["anchor_pos"] >= 0
    first_use = np.full(r1["anchor_count"], len(am), np.int64)
    idx = np.nonzero(am)[0]
    np.minimum.at(first_use, r1["anchor_addr"][am], idx)
    assert np.all(np.diff(first_use) > 0)
    eng.close()


@pytest.mark.parametrize("k,fastq,lossless,seed", [(31, True, False, 11), (63, True, True, 12), (31, False, False, 13)])
def test_indels_n_runs_variable_lengths(k, fastq, lossless, seed):
    """BASELINE configs[3]/[4] shape at reduced scale: 150 bp reads with indels (variable lengths), N runs, FASTQ lossy
    (k = 31) and lossless (k = 63), several blocks, vs the oracle."""
    from oracle import oracle_py as O
    from leon_b200 import synth
    g = synth.make_genome(50000, seed)
    bases, offs, quals = synth.sample_reads(g, 15000, 150, seed + 100, n_rate=0.001, with_quals=fastq)
    bases, offs, quals = synth.add_indels_and_n_runs(bases, offs, quals, seed + 200, indel_rate=0.002, n_read_rate=0.02)
    assert len(np.unique(np.diff(offs.astype(np.int64)))) > 3
    ref = O.compress(bases, offs, quals, kmer_size=k, reads_per_block=4000, lossless=lossless)
    eng = _engine(kmer_size=k, reads_per_block=4000, has_qual=fastq, lossless=lossless)
    r = eng.compress(bases, offs, quals)
    eng.close()
    _same(r, ref, fastq)



def test_configs3_shape_at_one_hundredth_scale():
    """BASELINE configs[3] at x0.01: 1 Mbp genome, 333 333 FASTQ reads of 150 bp with indels, N bases, N runs and
    trimmed reads, lossy, 50000-read blocks (7 blocks, 2 levels of the hash partition, several deflate chunks per block),
    byte for byte against the oracle."""
    from oracle import oracle_py as O
    from leon_b200 import workloads as W
    bases, offs, quals = W.generate_c3(1_000_000, 333_333)
    ref = O.compress(bases, offs, quals, kmer_size=31, reads_per_block=50000)
    eng = _engine(kmer_size=31, reads_per_block=50000, has_qual=True)
    r = eng.compress(bases, offs, quals)
    eng.close()
    _same(r, ref, True)
    # the same set with the quality blocks deflated on the device: every stream inflates to the reference's text
    import zlib
    eng = _engine(kmer_size=31, reads_per_block=50000, has_qual=True, qual_zlib_on_device=True, skip_qual_text=True, skip_read_arrays=True)
    r = eng.compress(bases, offs, quals)
    eng.close()
    assert len(r["anchor_pos"]) == 0 and len(r["qual_text"]) == 0
    at, parts = 0, []
    for nz in r["qual_z_block_sizes"]:
        parts.append(zlib.decompress(r["qual_z"][at:at + int(nz)].tobytes()))
        at += int(nz)
    assert b"".join(parts) == ref["qual_text"].tobytes()
    assert np.array_equal(r["dna"], ref["dna"]) and np.array_equal(r["anchors_dict"], ref["anchors_dict"])


def _long_reads(seed, genome_len, lengths, n_rate=0.001, fastq=False):
    """Reads of the given lengths sampled from a random genome with 0.5 % substitutions (both strands)."""
    from leon_b200 import synth
    rng = np.random.Generator(np.random.PCG64(seed))
    g = synth.make_genome(genome_len, seed)
    bs, qs = [], []
    for L in lengths:
        p = int(rng.integers(0, genome_len - L + 1))
        r = g[p:p + L].copy()
        if rng.random() < 0.5:
            r = synth._COMP[r[::-1]]
        e = rng.random(L) < 0.005
        r[e] = synth.ACGT[(np.searchsorted(synth.ACGT, r[e]) + rng.integers(1, 4, int(e.sum()))) % 4]
        r[rng.random(L) < n_rate] = ord("N")
        bs.append(r)
        if fastq:
            q = rng.integers(2, 41, L).astype(np.uint8) + 33
            qs.append(q)
    offs = np.zeros(len(lengths) + 1, np.uint64)
    np.cumsum(np.array(lengths, np.int64), out=offs[1:])
    return np.concatenate(bs), offs, (np.concatenate(qs) if fastq else None)


@pytest.mark.parametrize("k,fastq,kind,seed", [(31, False, "miseq250", 31), (31, True, "miseq300", 32), (63, False, "kb", 33),
                                               (31, True, "mixed", 34), (63, True, "mixed", 35), (31, False, "kb", 36)])
def test_long_reads_unstaged_paths(k, fastq, kind, seed):
    """Reads longer than the shared-memory staging of k_extract / k_anchor_lookup / k_walk (159 bp) and of k_quals: 250 /
    300 bp, several kb, and short and long reads mixed inside one CTA tile -- the global-memory paths, byte for byte."""
    from oracle import oracle_py as O
    rng = np.random.Generator(np.random.PCG64(seed))
    if kind == "miseq250":
        lengths = [250] * 6000
    elif kind == "miseq300":
        lengths = [int(x) for x in rng.integers(200, 301, 5000)]
    elif kind == "kb":
        lengths = [int(x) for x in rng.integers(1500, 9000, 400)]
    else:
        lengths = [int(x) for x in rng.choice([40, 100, 151, 250, 1000, 4000], 1500, p=[0.1, 0.4, 0.2, 0.2, 0.07, 0.03])]
    bases, offs, quals = _long_reads(seed, 60000, lengths, fastq=fastq)
    rpb = 700
    ref = O.compress(bases, offs, quals, kmer_size=k, reads_per_block=rpb)
    eng = _engine(kmer_size=k, reads_per_block=rpb, has_qual=fastq)
    r = eng.compress(bases, offs, quals)
    eng.close()
    _same(r, ref, fastq)


def test_create_destroy_does_not_leak():
    """leon_destroy returns every device and pinned allocation of the context (FASTQ path with device deflate included)."""
    import torch
    from leon_b200 import synth
    g = synth.make_genome(30000, 5)
    bases, offs, quals = synth.sample_reads(g, 20000, 100, 6, with_quals=True)
    free = []
    for i in range(4):
        eng = _engine(kmer_size=31, reads_per_block=5000, has_qual=True, qual_zlib_on_device=True)
        eng.compress(bases, offs, quals)
        eng.close()
        torch.cuda.synchronize()
        free.append(torch.cuda.mem_get_info()[0])
    assert free[-1] >= free[1] - (8 << 20), free      # after the first cycle (context warm-up) nothing is held back


@pytest.mark.parametrize("k,fastq,lossless,passes,batch_blocks,seed", [(31, True, False, 1, 3, 41), (31, False, False, 3, 2, 42),
                                                                      (63, True, True, 2, 5, 43), (31, True, False, 4, 1, 44)])
def test_streaming_equals_one_shot(k, fastq, lossless, passes, batch_blocks, seed):
    """Streaming entry points (leon_count_begin / leon_count / leon_count_finish / leon_encode_block / leon_encode_finish):
    the set fed in batches of a few blocks, counted in 1-4 hash-partitioned passes, gives the bytes of the oracle (and so of
    the one-shot path), whatever the batching; configs[4] shape (k = 63, lossless) included."""
    from oracle import oracle_py as O
    from leon_b200 import workloads as W
    rpb = 3000
    bases, offs, quals = W.generate_c3(80_000, 26_500, gseed=seed, rseed=seed + 1, with_quals=fastq)
    ref = O.compress(bases, offs, quals, kmer_size=k, reads_per_block=rpb, lossless=lossless)
    eng = _engine(kmer_size=k, reads_per_block=rpb, has_qual=fastq, lossless=lossless)
    r = eng.compress_stream(bases, offs, quals, batch_reads=batch_blocks * rpb, count_passes=passes)
    assert r["count_passes"] == passes
    _same(r, ref, fastq)
    # a second set on the same context, other batching, passes chosen by the library (1 at this size); every next batch
    # announced before the current one is submitted (two announcements open at a time: include/leon_b200.h, leon_stream_prefetch)
    r2 = eng.compress_stream(bases, offs, quals, batch_reads=4 * rpb, count_passes=0, prefetch="all")
    assert r2["count_passes"] == 1
    _same(r2, ref, fastq)
    r3 = eng.compress_stream(bases, offs, quals, batch_reads=2 * rpb, count_passes=2, prefetch="all" if seed % 2 else "none")
    eng.close()
    _same(r3, ref, fastq)


def test_streaming_device_memory_is_independent_of_the_number_of_reads():
    """The device footprint of the streaming path is set by the batch size, not by the number of reads: 4x the reads in the
    same batches must not need more device memory (beyond the solid k-mer list and the dictionary, which follow the genome)."""
    import torch
    from leon_b200 import workloads as W
    rpb = 5000
    used = []
    for nreads in (60_000, 240_000):
        bases, offs, quals = W.generate_c3(200_000, nreads, with_quals=True)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        eng = _engine(kmer_size=31, reads_per_block=rpb, has_qual=True, qual_zlib_on_device=True, skip_qual_text=True, skip_read_arrays=True)
        r = eng.compress_stream(bases, offs, quals, batch_reads=4 * rpb, count_passes=4 if nreads > 100_000 else 1)
        torch.cuda.synchronize()
        used.append(free0 - torch.cuda.mem_get_info()[0])
        assert r["read_count"] == nreads and int(r["blocksizes"][1::2].sum()) == nreads
        eng.close()
    assert used[1] < used[0] * 1.5 + (64 << 20), used


def test_pipelined_stage_many_superblocks(monkeypatch):
    """More super-blocks than SB_PIPELINE_MIN: the walk / entropy stage of a super-block runs beside the later anchor
    batches (the path large read sets take), byte for byte against the oracle."""
    from oracle import oracle_py as O
    from leon_b200 import synth
    g = synth.make_genome(40000, 77)
    bases, offs, quals = synth.sample_reads(g, 30000, 100, 78, n_rate=0.0005, with_quals=True)
    ref = O.compress(bases, offs, quals, kmer_size=31, reads_per_block=500)
    eng = _engine(kmer_size=31, reads_per_block=500, has_qual=True, anchor_batch=1000)     # 30 super-blocks of 2 blocks
    r = eng.compress(bases, offs, quals)
    eng.close()
    _same(r, ref, True)
