"""Host-side dealing of a read set to the ranks of the multi-GPU path, and reassembly of their outputs.

Rule (include/leon_b200.h, leon_comm_*): the reads are dealt in rounds of nranks * S consecutive reads, S to each rank
in rank order; S = Leon.comm_batch_reads() is a multiple of reads_per_block, so blocks stay whole and local block j
of round g on rank r is block (g * nranks + r) * S / reads_per_block + j of the file.  numpy only, no compute.
"""
import numpy as np


def rank_read_ranges(nreads, rank, nranks, S):
    """[(first, last)) ranges of the global read indices that belong to `rank`, in file order."""
    out = []
    g = 0
    while True:
        lo = (g * nranks + rank) * S
        if lo >= nreads:
            break
        out.append((lo, min(nreads, lo + S)))
        g += 1
    return out


def shard_reads(bases, offsets, quals, rank, nranks, S):
    """The reads of `rank`: (bases, offsets, quals) in the layout leon_upload expects."""
    offsets = np.asarray(offsets, np.uint64)
    n = len(offsets) - 1
    ranges = rank_read_ranges(n, rank, nranks, S)
    if not ranges:
        return np.zeros(0, np.uint8), np.zeros(1, np.uint64), (np.zeros(0, np.uint8) if quals is not None else None)
    parts_b, parts_q, lens = [], [], []
    for lo, hi in ranges:
        b0, b1 = int(offsets[lo]), int(offsets[hi])
        parts_b.append(bases[b0:b1])
        if quals is not None:
            parts_q.append(quals[b0:b1])
        lens.append(np.diff(offsets[lo:hi + 1]))
    L = np.concatenate(lens)
    offs = np.zeros(len(L) + 1, np.uint64)
    np.cumsum(L, out=offs[1:])
    return (np.concatenate(parts_b), offs, np.concatenate(parts_q) if quals is not None else None)


def global_block_ids(nreads, rank, nranks, S, reads_per_block):
    """File block id of every local block of `rank`, in local order."""
    assert S % reads_per_block == 0
    ids = []
    for lo, hi in rank_read_ranges(nreads, rank, nranks, S):
        first = lo // reads_per_block
        ids.extend(range(first, first + (hi - lo + reads_per_block - 1) // reads_per_block))
    return ids


def stream_batches(nreads, rank, nranks, S, chunks_per_batch):
    """Streaming path on several ranks (leon_count / leon_encode_block, include/leon_b200.h): the (first, last) LOCAL read
    indices of the batches `rank` feeds, `chunks_per_batch` rounds of S reads at a time.  Every rank makes the same number of
    calls -- the anchor exchange of a round is a collective -- so a rank that has run out of reads goes on with empty batches."""
    local = sum(hi - lo for lo, hi in rank_read_ranges(nreads, rank, nranks, S))
    nchunks = (nreads + S - 1) // S
    max_chunks = (nchunks + nranks - 1) // nranks
    M = max(1, int(chunks_per_batch))
    ncalls = max(1, (max_chunks + M - 1) // M)
    return [(min(local, i * M * S), min(local, (i + 1) * M * S)) for i in range(ncalls)]


def merge_results(results, nreads, nranks, S, reads_per_block):
    """Reassembles the per-rank result dicts (Leon._result) into the single-GPU layout: blocks in file order, per-read
    arrays in file order; whole-set fields (Bloom, counts, anchor dictionary) come from rank 0."""
    nblocks = (nreads + reads_per_block - 1) // reads_per_block
    dna_parts = [None] * nblocks
    q_parts = [None] * nblocks
    bs = np.zeros(2 * nblocks, np.uint64)
    qs = np.zeros(nblocks, np.uint64)
    has_q = any(len(r["qual_text"]) for r in results)
    has_z = any(len(r.get("qual_z", ())) for r in results)
    z_parts = [None] * nblocks
    zs = np.zeros(nblocks, np.uint64)
    apos = np.zeros(nreads, np.int32)
    aaddr = np.zeros(nreads, np.uint32)
    for rank, r in enumerate(results):
        ids = global_block_ids(nreads, rank, nranks, S, reads_per_block)
        assert len(ids) == r["nblocks"], (rank, len(ids), r["nblocks"])
        off = qoff = zoff = 0
        for j, b in enumerate(ids):
            nb = int(r["blocksizes"][2 * j])
            dna_parts[b] = r["dna"][off:off + nb]
            off += nb
            bs[2 * b:2 * b + 2] = r["blocksizes"][2 * j:2 * j + 2]
            if has_q:
                nq = int(r["qual_block_sizes"][j])
                q_parts[b] = r["qual_text"][qoff:qoff + nq]
                qoff += nq
                qs[b] = nq
            if has_z:
                nz = int(r["qual_z_block_sizes"][j])
                z_parts[b] = r["qual_z"][zoff:zoff + nz]
                zoff += nz
                zs[b] = nz
        loc = 0
        for lo, hi in rank_read_ranges(nreads, rank, nranks, S):
            if not len(r["anchor_pos"]):     # skip_read_arrays: the per-read parity arrays were not fetched
                break
            apos[lo:hi] = r["anchor_pos"][loc:loc + hi - lo]
            aaddr[lo:hi] = r["anchor_addr"][loc:loc + hi - lo]
            loc += hi - lo
    out = dict(results[0])
    out.update(
        nblocks=nblocks, blocksizes=bs,
        dna=np.concatenate(dna_parts) if nblocks else np.zeros(0, np.uint8),
        qual_text=np.concatenate(q_parts) if (has_q and nblocks) else np.zeros(0, np.uint8),
        qual_block_sizes=qs if has_q else np.zeros(0, np.uint64),
        qual_z=np.concatenate(z_parts) if (has_z and nblocks) else np.zeros(0, np.uint8),
        qual_z_block_sizes=zs if has_z else np.zeros(0, np.uint64),
        anchor_pos=apos, anchor_addr=aaddr,
        read_count=sum(r["read_count"] for r in results), total_dna_size=sum(r["total_dna_size"] for r in results),
        min_read_size=min(r["min_read_size"] for r in results if r["read_count"]) if nreads else results[0]["min_read_size"],
        max_read_size=max(r["max_read_size"] for r in results),
        reads_without_anchor=sum(r["reads_without_anchor"] for r in results),
    )
    return out


def choose_batch_reads(nreads, nranks, reads_per_block, max_blocks=16):
    """Reads per rank and anchor batch (S, a multiple of reads_per_block) that balances the ranks for a set of known
    size: the S in [1, max_blocks] blocks minimising (blocks of the busiest rank + number of exchange rounds) -- a round
    of anchor exchanges costs about as much as coding one block.  Any S gives the same output; this only shortens the run."""
    nblocks = (nreads + reads_per_block - 1) // reads_per_block
    best = None
    for sb in range(1, max_blocks + 1):
        per = [0] * nranks
        rounds = 0
        b = 0
        while b < nblocks:
            for r in range(nranks):
                take = min(sb, max(0, nblocks - b))
                per[r] += take
                b += take
            rounds += 1
        key = max(per) + rounds
        if best is None or key <= best[0]:
            best = (key, sb)
    return best[1] * reads_per_block
