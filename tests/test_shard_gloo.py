"""Host side of the multi-GPU path on CPU: two gloo ranks deal a seeded read set (leon_b200.shard), exchange per-rank
results shaped like Leon._result and merge them back; the merged layout must equal the single-set one.  The per-rank
"results" are cut out of the ORACLE's output for the whole set (test infrastructure; the CUDA path itself is covered by
tools/multigpu_check.py on real GPUs), so this checks exactly the dealing, block numbering and merging logic."""
import os
import socket

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_result(full, bases, offs, quals, rank, world, S, rpb):
    """What rank `rank` would return: its own blocks / reads, cut out of the whole-set result."""
    from leon_b200 import shard
    n = len(offs) - 1
    ids = shard.global_block_ids(n, rank, world, S, rpb)
    bs = full["blocksizes"]
    starts = np.concatenate([[0], np.cumsum(bs[0::2])]).astype(np.int64)
    dna = [full["dna"][starts[b]:starts[b + 1]] for b in ids]
    has_q = len(full["qual_block_sizes"]) > 0
    if has_q:
        qstarts = np.concatenate([[0], np.cumsum(full["qual_block_sizes"])]).astype(np.int64)
        qt = [full["qual_text"][qstarts[b]:qstarts[b + 1]] for b in ids]
    ranges = shard.rank_read_ranges(n, rank, world, S)
    sb, so, sq = shard.shard_reads(bases, offs, quals, rank, world, S)
    lens = np.diff(so.astype(np.int64))
    r = dict(full)
    r.update(
        nblocks=len(ids),
        blocksizes=np.array([v for b in ids for v in (bs[2 * b], bs[2 * b + 1])], np.uint64),
        dna=np.concatenate(dna) if dna else np.zeros(0, np.uint8),
        qual_text=np.concatenate(qt) if (has_q and ids) else np.zeros(0, np.uint8),
        qual_block_sizes=np.array([full["qual_block_sizes"][b] for b in ids], np.uint64) if has_q else np.zeros(0, np.uint64),
        anchor_pos=np.concatenate([full["anchor_pos"][lo:hi] for lo, hi in ranges]) if ranges else np.zeros(0, np.int32),
        anchor_addr=np.concatenate([full["anchor_addr"][lo:hi] for lo, hi in ranges]) if ranges else np.zeros(0, np.uint32),
        read_count=len(so) - 1, total_dna_size=int(so[-1]),
        min_read_size=int(lens.min()) if len(lens) else 0, max_read_size=int(lens.max()) if len(lens) else 0,
        reads_without_anchor=int(sum((full["anchor_pos"][lo:hi] < 0).sum() for lo, hi in ranges)),
    )
    assert len(sb) == int(so[-1])
    return r


def _worker(rank, world, port, nreads, rpb, fastq, S, out_path):
    import sys
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from leon_b200 import shard, synth
    from oracle import oracle_py as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = synth.make_genome(8000, 7)
    bases, offs, quals = synth.sample_reads(g, nreads, 60, 8, n_rate=0.001, with_quals=fastq)
    # ragged lengths: drop a few bases from some reads
    full = O.compress(bases, offs, quals, kmer_size=21, reads_per_block=rpb)
    mine = _rank_result(full, bases, offs, quals, rank, world, S, rpb)
    allr = [None] * world
    dist.all_gather_object(allr, mine)
    ok = True
    if rank == 0:
        m = shard.merge_results(allr, nreads, world, S, rpb)
        for key in ("dna", "blocksizes", "anchor_pos", "anchor_addr", "qual_text", "qual_block_sizes"):
            ok &= bool(np.array_equal(m[key], full[key]))
        for key in ("nblocks", "read_count", "total_dna_size", "min_read_size", "max_read_size", "reads_without_anchor"):
            ok &= m[key] == full[key]
        # every read belongs to exactly one rank, in file order inside a rank
        seen = np.zeros(nreads, np.int32)
        for r in range(world):
            for lo, hi in shard.rank_read_ranges(nreads, r, world, S):
                seen[lo:hi] += 1
        ok &= bool((seen == 1).all())
        open(out_path, "w").write("OK" if ok else "FAILED")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("nreads,rpb,fastq,S", [(2300, 100, False, 300), (1750, 250, True, 250), (90, 100, False, 200)])
def test_two_gloo_ranks_deal_and_merge(tmp_path, nreads, rpb, fastq, S):
    import torch.multiprocessing as mp
    from oracle import oracle_py as O
    O.build()
    out = str(tmp_path / "result.txt")
    mp.spawn(_worker, args=(2, _free_port(), nreads, rpb, fastq, S, out), nprocs=2, join=True)
    assert open(out).read() == "OK"


def test_choose_batch_reads_balances():
    from leon_b200 import shard
    for n in (2, 4, 8):
        S = shard.choose_batch_reads(4_600_000, n, 50000)
        assert S % 50000 == 0
        per = [sum(hi - lo for lo, hi in shard.rank_read_ranges(4_600_000, r, n, S)) for r in range(n)]
        assert sum(per) == 4_600_000
        assert max(per) <= 1.35 * 4_600_000 / n


def _stream_worker(rank, world, port, nreads, S, rpb, M, out_path):
    import sys
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from leon_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    batches = shard.stream_batches(nreads, rank, world, S, M)
    ranges = shard.rank_read_ranges(nreads, rank, world, S)
    local = sum(hi - lo for lo, hi in ranges)
    ok = batches[0][0] == 0 and batches[-1][1] == local and all(a[1] == b[0] for a, b in zip(batches, batches[1:]))
    # every rank makes the same number of calls (the anchor exchange of a round is a collective)
    t = torch.tensor([len(batches)])
    lo_t, hi_t = t.clone(), t.clone()
    dist.all_reduce(lo_t, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi_t, op=dist.ReduceOp.MAX)
    ok &= int(lo_t) == int(hi_t)
    # the engine numbers the reads of local round i of rank r from (i * world + r) * S (leon_b200.cu encode_batch, delta0):
    # walking the batches round by round must give back exactly this rank's global ranges
    seen = []
    done_rounds = 0
    for lo, hi in batches:
        at = lo
        while at < hi:
            n = min(S, hi - at)
            g0 = (done_rounds * world + rank) * S
            seen.append((g0, g0 + n))
            at += n
            done_rounds += 1
    ok &= seen == ranges
    mine = dict(ids=shard.global_block_ids(nreads, rank, world, S, rpb), ok=bool(ok))
    allr = [None] * world
    dist.all_gather_object(allr, mine)
    if rank == 0:
        ids = sorted(i for r in allr for i in r["ids"])
        good = all(r["ok"] for r in allr) and ids == list(range((nreads + rpb - 1) // rpb))
        open(out_path, "w").write("OK" if good else "FAILED")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("nreads,S,rpb,M", [(2300, 300, 100, 2), (1750, 250, 50, 1), (90, 200, 100, 4), (4000, 500, 100, 3)])
def test_two_gloo_ranks_streaming_batches(tmp_path, nreads, S, rpb, M):
    import torch.multiprocessing as mp
    out = str(tmp_path / "result.txt")
    mp.spawn(_stream_worker, args=(2, _free_port(), nreads, S, rpb, M, out), nprocs=2, join=True)
    assert open(out).read() == "OK"
