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
