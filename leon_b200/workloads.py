"""The BASELINE.json workloads as seeded generators, the file writer the reference arm uses, and the output digest.

One place defines what "configs[1]" / "configs[2]" / "configs[3]" are, so that bench.py, the digest generator
(tools/make_digest.py, which runs the reference with -nb-cores 1) and the tests all see the same bytes.  numpy only.
"""
import hashlib
import zlib

import numpy as np

from . import synth

KMER = 31
RPB = 50000

# name -> (genome length, genome seed, reads, read length, read seed, fastq, n_rate, description)
CONFIGS = {
    "configs[1]": dict(genome=4_600_000, gseed=1, reads=4_600_000, read_len=100, rseed=2, fastq=False, n_rate=0.0,
                       text="configs[1]: 4.6 Mbp random genome, 100 bp FASTA reads at 100x, 0.5% substitutions, k=31, seq-only",
                       ref_flags=["-seq-only"]),
    "configs[2]": dict(genome=4_600_000, gseed=1, reads=4_600_000, read_len=100, rseed=2, fastq=True, n_rate=0.0,
                       text="configs[2]: 4.6 Mbp random genome, 100 bp FASTQ reads at 100x, 0.5% substitutions, Illumina-like "
                            "synthetic qualities, lossy quality mode, k=31 (sequence + quality streams; names not kept)",
                       ref_flags=["-noheader"]),
}


def generate(name, out_bases=None, out_quals=None, reads=None):
    """(bases, offsets, quals) of a named workload; `reads` overrides the read count (debug sizes)."""
    c = CONFIGS[name]
    g = synth.make_genome(c["genome"], c["gseed"])
    return synth.sample_reads(g, reads if reads else c["reads"], c["read_len"], c["rseed"], n_rate=c["n_rate"],
                              with_quals=c["fastq"], out_bases=out_bases, out_quals=out_quals)


def write_fixed(path, bases, read_len, quals=None, chunk=1 << 18):
    """FASTA / FASTQ writer for fixed-length reads, names r<i> (same text as synth.write_fasta, vectorised)."""
    n = len(bases) // read_len
    with open(path, "wb") as f:
        lo = 0
        while lo < n:
            hi = min(n, lo + chunk)
            # split at powers of ten so that every row of a piece has the same number of digits
            nd = len(str(lo))
            hi = min(hi, 10 ** nd)
            m = hi - lo
            idx = np.arange(lo, hi, dtype=np.int64)
            w = 2 + nd + 1 + read_len + 1 + (0 if quals is None else 2 + read_len + 1)
            rec = np.empty((m, w), np.uint8)
            rec[:, 0] = ord(">") if quals is None else ord("@")
            rec[:, 1] = ord("r")
            for d in range(nd):
                rec[:, 2 + d] = (idx // 10 ** (nd - 1 - d)) % 10 + ord("0")
            p = 2 + nd
            rec[:, p] = 10
            rec[:, p + 1:p + 1 + read_len] = bases[lo * read_len:hi * read_len].reshape(m, read_len)
            p += 1 + read_len
            rec[:, p] = 10
            if quals is not None:
                rec[:, p + 1] = ord("+")
                rec[:, p + 2] = 10
                rec[:, p + 3:p + 3 + read_len] = quals[lo * read_len:hi * read_len].reshape(m, read_len)
                rec[:, p + 3 + read_len] = 10
            f.write(rec.tobytes())
            lo = hi


def _sha(*parts):
    h = hashlib.sha256()
    for p in parts:
        h.update(memoryview(np.ascontiguousarray(p)).cast("B"))
    return h.hexdigest()


def block_digests(blocks):
    """sha256 of every block (an iterable of bytes-like objects), as a list of 32-byte digests."""
    return [hashlib.sha256(memoryview(np.ascontiguousarray(b)).cast("B") if isinstance(b, np.ndarray) else b).digest() for b in blocks]


def digest_of_streams(bloom, anchors_dict, dna, blocksizes, qual_text_blocks=None):
    """Digest of everything `leon -c` stores for the sequence (+ quality) streams.  qual_text_blocks: iterable of the
    INFLATED quality text of every block in order (zlib bytes may differ between writers, the text may not).
    dna_cc / qual_text_cc are checksums of the per-block checksums (sha256 over the blocks' sha256 digests in file order): ranks
    of a multi-GPU run can each hash their own blocks and only the 32-byte digests need to travel."""
    bs = np.ascontiguousarray(blocksizes, "<u8")
    dna = np.ascontiguousarray(dna)
    ends = np.cumsum(bs[0::2].astype(np.int64))
    starts = ends - bs[0::2].astype(np.int64)
    d = {"bloom": _sha(bloom), "anchors_dict": _sha(anchors_dict), "dna": _sha(dna), "blocksizes": _sha(bs),
         "nblocks": int(len(bs) // 2), "dna_nbytes": int(len(dna)),
         "dna_cc": hashlib.sha256(b"".join(block_digests(dna[int(a):int(b)] for a, b in zip(starts, ends)))).hexdigest()}
    if qual_text_blocks is not None:
        h = hashlib.sha256()
        n = 0
        per = []
        for t in qual_text_blocks:
            h.update(t)
            per.append(hashlib.sha256(t).digest())
            n += len(t)
        d["qual_text"] = h.hexdigest()
        d["qual_text_nbytes"] = n
        d["qual_text_cc"] = hashlib.sha256(b"".join(per)).hexdigest()
    return d


def digest_of_result(r, with_qual):
    """Digest of a result dict of leon_b200.Leon (single GPU, or shard.merge_results of all ranks)."""
    qt = None
    if with_qual:
        if len(r.get("qual_z", ())):
            sizes = np.asarray(r["qual_z_block_sizes"], np.uint64)
            z = memoryview(np.ascontiguousarray(r["qual_z"]))
            offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
            qt = (zlib.decompress(z[int(offs[b]):int(offs[b + 1])]) for b in range(len(sizes)))
        else:
            sizes = np.asarray(r["qual_block_sizes"], np.uint64)
            t = memoryview(np.ascontiguousarray(r["qual_text"]))
            offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
            qt = (t[int(offs[b]):int(offs[b + 1])] for b in range(len(sizes)))
    return digest_of_streams(r["bloom"], r["anchors_dict"], r["dna"], r["blocksizes"], qt)


def digest_of_container(path, with_qual):
    """Digest of a `.leon` container written by the reference (or by us), read with the library's own reader."""
    from . import api
    with api.Container(path) as c:
        bs = np.frombuffer(c.dataset("leon/dna/blocksizes").tobytes(), "<u8")
        nblocks = len(bs) // 2
        bs = bs[:2 * nblocks]
        dna = [c.dataset("leon/dna/dna_%d" % b)[:int(bs[2 * b])] for b in range(nblocks) if bs[2 * b]]
        dna = np.concatenate(dna) if dna else np.zeros(0, np.uint8)
        qt = None
        if with_qual:
            qt = (zlib.decompressobj().decompress(c.dataset("leon/qual/qual_%d" % b).tobytes()) for b in range(nblocks))
        return digest_of_streams(c.dataset("leon/bloom"), c.dataset("leon/anchors/anchorsDict"), dna, bs, qt)


def write_ragged(path, bases, offsets, quals, first_index=0, append=False, chunk=1 << 16):
    """FASTA / FASTQ writer for reads of any length, names r<i>."""
    n = len(offsets) - 1
    with open(path, "ab" if append else "wb") as f:
        for lo in range(0, n, chunk):
            hi = min(n, lo + chunk)
            parts = []
            for i in range(lo, hi):
                a, b = int(offsets[i]), int(offsets[i + 1])
                if quals is None:
                    parts.append(b">r%d\n%s\n" % (first_index + i, bases[a:b].tobytes()))
                else:
                    parts.append(b"@r%d\n%s\n+\n%s\n" % (first_index + i, bases[a:b].tobytes(), quals[a:b].tobytes()))
            f.write(b"".join(parts))


def digests_equal(a, b):
    keys = ["bloom", "anchors_dict", "dna", "blocksizes"] + (["qual_text"] if "qual_text" in a and "qual_text" in b else [])
    return all(a.get(k) == b.get(k) for k in keys), [k for k in keys if a.get(k) != b.get(k)]


# ---------------------------------------------------------------------------------------------------------------
# configs[3] shape: 150 bp FASTQ at 50x with N bases, N runs and indels (fixed-length reads whose alignment to the genome
# has an insertion or a deletion), a few reads trimmed (variable lengths).  Generated in chunks of 250 000 reads, chunk c
# from its own generator PCG64([seed, c]), so that any rank can produce exactly its own reads and threads can work on
# different chunks at once.
# ---------------------------------------------------------------------------------------------------------------
C3_CHUNK = 250_000        # reads; 5 blocks of 50000, so a chunk is also a unit the reads can be dealt to ranks in
CONFIGS["configs[3]"] = dict(genome=100_000_000, gseed=3, reads=33_333_333, read_len=150, rseed=4, fastq=True, n_rate=0.001,
                             text="configs[3]: 100 Mbp random genome, 150 bp FASTQ reads at 50x, 0.5% substitutions, N bases and "
                                  "N runs, indels, 2% trimmed reads, lossy quality mode, k=31 (Bloom filter 150 MB > L2)",
                             ref_flags=["-noheader"])


def c3_chunk(genome, c, n_reads, read_len, seed, with_quals=True, sub_rate=0.005, n_rate=0.001, indel_rate=0.05,
             nrun_rate=0.005, trim_rate=0.02):
    """Reads [c * C3_CHUNK, min(n_reads, (c + 1) * C3_CHUNK)) of the configs[3]-shaped set: (bases, lengths, quals)."""
    lo, hi = c * C3_CHUNK, min(n_reads, (c + 1) * C3_CHUNK)
    m = hi - lo
    rng = np.random.Generator(np.random.PCG64([seed, c]))
    G = len(genome)
    L = read_len
    ar = np.arange(L, dtype=np.int64)[None, :]
    pos = rng.integers(0, G - L - 4, m)[:, None]
    # indel: kind 0 none, 1 deletion (m_ genome bases skipped at p), 2 insertion (m_ random bases at p)
    kind = np.where(rng.random(m) < indel_rate, rng.integers(1, 3, m), 0)[:, None]
    p = rng.integers(5, L - 5, m)[:, None]
    ml = rng.integers(1, 4, m)[:, None]
    idx = pos + ar + np.where((kind == 1) & (ar >= p), ml, 0) - np.where((kind == 2) & (ar >= p + ml), ml, 0)
    r = genome[idx]
    ins = (kind == 2) & (ar >= p) & (ar < p + ml)
    ni = int(ins.sum())
    if ni:
        r[ins] = ACGT_[rng.integers(0, 4, ni)]
    rev = rng.random(m) < 0.5
    r[rev] = synth._COMP[r[rev][:, ::-1]]
    err = rng.random((m, L)) < sub_rate
    ne = int(err.sum())
    if ne:
        cur = np.searchsorted(ACGT_, r[err])
        r[err] = ACGT_[(cur + rng.integers(1, 4, ne)) % 4]
    nmask = rng.random((m, L)) < n_rate
    has_run = (rng.random(m) < nrun_rate)[:, None]
    rp = rng.integers(0, L - 30, m)[:, None]
    rl = rng.integers(5, 31, m)[:, None]
    nmask |= has_run & (ar >= rp) & (ar < rp + rl)
    r[nmask] = ord("N")
    lens = np.where(rng.random(m) < trim_rate, rng.integers(20, L, m), L)
    keep = ar < lens[:, None]
    q = None
    if with_quals:
        qtab = np.array([30, 32, 35, 37, 38, 40, 40, 40], np.uint8)
        q = qtab[rng.integers(0, len(qtab), (m, L))]
        tail = L - 20
        dec = rng.integers(0, 11, (m, 20)).astype(np.int16)
        q[:, tail:] = np.maximum(2, q[:, tail:].astype(np.int16) - dec).astype(np.uint8)
        if ne:
            q[err] = rng.integers(2, 16, ne).astype(np.uint8)
        q[nmask] = 2
        q = (q + 33)[keep]
    return r[keep], lens.astype(np.int64), q


ACGT_ = synth.ACGT


def generate_c3(genome_len, n_reads, read_len=150, gseed=3, rseed=4, with_quals=True, chunks=None, threads=0):
    """configs[3]-shaped read set (or only the given chunk ids, in that order): (bases, offsets, quals)."""
    import concurrent.futures as cf
    import os
    g = synth.make_genome(genome_len, gseed)
    ids = list(range((n_reads + C3_CHUNK - 1) // C3_CHUNK)) if chunks is None else list(chunks)
    threads = threads or min(16, os.cpu_count() or 1)
    with cf.ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(lambda c: c3_chunk(g, c, n_reads, read_len, rseed, with_quals), ids))
    lens = np.concatenate([p[1] for p in parts]) if parts else np.zeros(0, np.int64)
    offs = np.zeros(len(lens) + 1, np.uint64)
    np.cumsum(lens, out=offs[1:])
    bases = np.concatenate([p[0] for p in parts]) if parts else np.zeros(0, np.uint8)
    quals = np.concatenate([p[2] for p in parts]) if (parts and with_quals) else None
    return bases, offs, quals
