"""Seeded synthetic read sets of the BASELINE.json shapes: reads sampled from a uniform random genome, both strands,
substitution errors, optional N bases and Illumina-like qualities.  numpy only (vectorised, chunked)."""
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, np.uint8)
for a, b in zip(b"ACGTN", b"TGCAN"):
    _COMP[a] = b


def make_genome(length, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    return ACGT[rng.integers(0, 4, length, dtype=np.uint8)]


def sample_reads(genome, n_reads, read_len, seed, sub_rate=0.005, n_rate=0.0, with_quals=False, out_bases=None,
                 out_quals=None, chunk=1 << 18):
    """Fixed-length reads.  Returns (bases u8[n_reads*read_len], offsets u64[n_reads+1], quals or None).
    out_bases / out_quals: optional preallocated (e.g. pinned) arrays to fill in place."""
    rng = np.random.Generator(np.random.PCG64(seed))
    G = len(genome)
    total = n_reads * read_len
    bases = out_bases if out_bases is not None else np.empty(total, np.uint8)
    quals = None
    if with_quals:
        quals = out_quals if out_quals is not None else np.empty(total, np.uint8)
    ar = np.arange(read_len, dtype=np.int64)
    qtab = np.array([30, 32, 35, 37, 38, 40, 40, 40], np.uint8)
    for lo in range(0, n_reads, chunk):
        hi = min(n_reads, lo + chunk)
        m = hi - lo
        pos = rng.integers(0, G - read_len + 1, m)
        r = genome[pos[:, None] + ar[None, :]]
        rev = rng.random(m) < 0.5
        r[rev] = _COMP[r[rev][:, ::-1]]
        err = rng.random((m, read_len)) < sub_rate
        ne = int(err.sum())
        if ne:
            cur = np.searchsorted(ACGT, r[err])           # ACGT is sorted in ASCII: A C G T
            r[err] = ACGT[(cur + rng.integers(1, 4, ne)) % 4]
        if n_rate:
            r[rng.random((m, read_len)) < n_rate] = ord("N")
        bases[lo * read_len:hi * read_len] = r.reshape(-1)
        if with_quals:
            q = qtab[rng.integers(0, len(qtab), (m, read_len))]
            tail = max(0, read_len - 20)
            dec = rng.integers(0, 11, (m, read_len - tail)).astype(np.int16)
            q[:, tail:] = np.maximum(2, q[:, tail:].astype(np.int16) - dec).astype(np.uint8)
            if ne:
                q[err] = rng.integers(2, 16, ne).astype(np.uint8)
            if n_rate:
                q[r == ord("N")] = 2
            quals[lo * read_len:hi * read_len] = (q + 33).reshape(-1)
    offsets = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(read_len)
    return bases, offsets, quals


def write_fasta(path, bases, offsets, quals=None):
    """Writes FASTA (or FASTQ when quals is given) with headers r<i>; used for the reference CPU baseline."""
    n = len(offsets) - 1
    with open(path, "wb") as f:
        for lo in range(0, n, 1 << 16):
            hi = min(n, lo + (1 << 16))
            parts = []
            for i in range(lo, hi):
                s = bases[int(offsets[i]):int(offsets[i + 1])].tobytes()
                if quals is None:
                    parts.append(b">r%d\n%s\n" % (i, s))
                else:
                    parts.append(b"@r%d\n%s\n+\n%s\n" % (i, s, quals[int(offsets[i]):int(offsets[i + 1])].tobytes()))
            f.write(b"".join(parts))


def add_indels_and_n_runs(bases, offsets, quals, seed, indel_rate=0.0005, n_read_rate=0.005):
    """BASELINE configs[3] shape: per-base insertions / deletions of 1-3 bases (so read lengths vary) and, in a fraction
    of the reads, one run of 5-30 N.  Returns new (bases, offsets, quals)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = len(offsets) - 1
    out_b, out_q, lens = [], [], np.zeros(n, np.int64)
    for i in range(n):
        lo, hi = int(offsets[i]), int(offsets[i + 1])
        r = bases[lo:hi].copy()
        q = quals[lo:hi].copy() if quals is not None else None
        L = len(r)
        nev = rng.binomial(L, indel_rate) if L else 0
        for _ in range(nev):
            p = int(rng.integers(0, max(1, len(r))))
            m = int(rng.integers(1, 4))
            if rng.random() < 0.5:                       # deletion
                r = np.concatenate([r[:p], r[p + m:]])
                if q is not None:
                    q = np.concatenate([q[:p], q[p + m:]])
            else:                                        # insertion
                ins = ACGT[rng.integers(0, 4, m)]
                r = np.concatenate([r[:p], ins, r[p:]])
                if q is not None:
                    q = np.concatenate([q[:p], np.full(m, 33 + 20, np.uint8), q[p:]])
        if len(r) and rng.random() < n_read_rate:
            m = int(rng.integers(5, 31))
            p = int(rng.integers(0, max(1, len(r) - m)))
            r[p:p + m] = ord("N")
            if q is not None:
                q[p:p + m] = 33 + 2
        out_b.append(r)
        if q is not None:
            out_q.append(q)
        lens[i] = len(r)
    offs = np.zeros(n + 1, np.uint64)
    np.cumsum(lens, out=offs[1:])
    return (np.concatenate(out_b) if out_b else np.zeros(0, np.uint8), offs,
            np.concatenate(out_q) if quals is not None else None)
