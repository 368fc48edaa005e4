#!/usr/bin/env python3
"""Parity of the multi-GPU path (run under torchrun, one rank per GPU): the merged output of N ranks must be byte
identical to the oracle (== the reference with -nb-cores 1) for seeded synthetic sets.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/multigpu_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import leon_b200  # noqa: E402
from leon_b200 import shard, synth  # noqa: E402

KEYS = ("cutoff", "nb_solids_cutoff", "nb_solid_total", "bloom_bits", "anchor_count", "read_count", "total_dna_size",
        "reads_without_anchor")
ARRS = ("histogram", "bloom", "anchors_dict", "blocksizes", "dna", "anchor_pos")


def run_case(rank, world, k, L, nreads, rpb, fastq, seed, genome, stream=0, passes=0, lossless=False):
    from oracle import oracle_py as O
    g = synth.make_genome(genome, seed)
    bases, offs, quals = synth.sample_reads(g, nreads, L, seed + 100, n_rate=0.0005, with_quals=fastq)
    eng = leon_b200.Leon(kmer_size=k, reads_per_block=rpb, has_qual=fastq, lossless=lossless, device=rank, qual_zlib_on_device=fastq)
    uid = [leon_b200.Leon.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    eng.comm_init(rank, world, uid[0])
    S = eng.comm_batch_reads()
    b, o, q = shard.shard_reads(bases, offs, quals, rank, world, S)
    if stream:      # streaming entry points: batches of `stream` rounds, the same number of calls on every rank
        rounds = -(-nreads // (world * S))
        nloc = len(o) - 1
        batches = [(min(nloc, i * stream * S), min(nloc, (i + 1) * stream * S)) for i in range(-(-rounds // stream))]
        r = eng.compress_stream(b, o, q, batches=batches, count_passes=passes)
    else:
        r = eng.compress(b, o, q)
    eng.close()
    allr = [None] * world
    dist.all_gather_object(allr, r)
    ok = True
    if rank == 0:
        m = shard.merge_results(allr, nreads, world, S, rpb)
        ref = O.compress(bases, offs, quals, kmer_size=k, reads_per_block=rpb, lossless=lossless)
        for key in KEYS:
            if m[key] != ref[key]:
                ok = False
                print("  MISMATCH %s: got %s want %s" % (key, m[key], ref[key]))
        for key in ARRS + (("qual_text", "qual_block_sizes") if fastq else ()):
            if not np.array_equal(m[key], ref[key]):
                ok = False
                d = np.nonzero(m[key][:min(len(m[key]), len(ref[key]))] != ref[key][:min(len(m[key]), len(ref[key]))])[0]
                print("  MISMATCH %s: len %d/%d first diff %s" % (key, len(m[key]), len(ref[key]), d[:1]))
        if fastq:       # quality blocks deflated on each rank's GPU: merged streams must inflate to the reference's text
            import zlib
            at, parts = 0, []
            for nz in m["qual_z_block_sizes"]:
                parts.append(zlib.decompress(m["qual_z"][at:at + int(nz)].tobytes()))
                at += int(nz)
            if b"".join(parts) != ref["qual_text"].tobytes() or [len(x) for x in parts] != [int(x) for x in ref["qual_block_sizes"]]:
                ok = False
                print("  MISMATCH inflated qual_z")
        am = ref["anchor_pos"] >= 0
        if not np.array_equal(m["anchor_addr"][am], ref["anchor_addr"][am]):
            ok = False
            print("  MISMATCH anchor_addr")
        print("case k=%d L=%d reads=%d rpb=%d fastq=%s world=%d S=%d stream=%d passes=%d: %s (anchors %d, dna %d bytes)" % (
            k, L, nreads, rpb, fastq, world, S, stream, passes, "OK" if ok else "FAILED", m["anchor_count"], len(m["dna"])))
    flag = [ok]
    dist.broadcast_object_list(flag, src=0)
    return flag[0]


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo")
    cases = [
        (31, 100, 120000, 2000, False, 1, 60000),
        (31, 100, 130500, 1000, True, 2, 40000),     # ragged: last round partial, some ranks short
        (63, 150, 60000, 1500, False, 3, 50000),
        (25, 76, 9000, 3000, False, 4, 20000),       # fewer rounds than ranks get reads
        (31, 100, 130500, 1000, True, 5, 40000, 3, 1),    # streaming: 3 rounds per call, one counting pass
        (31, 100, 120000, 2000, False, 6, 60000, 1, 3),   # streaming: one round per call, three counting passes
        (63, 150, 60000, 1500, True, 7, 50000, 2, 2),
        (63, 150, 90000, 1500, True, 8, 70000, 2, 3, True),   # configs[4] shape: k = 63, lossless, streamed, 3 counting passes
    ]
    ok = True
    for c in cases:
        ok &= run_case(rank, world, *c)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("ALL OK" if ok else "SOME FAILED")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
