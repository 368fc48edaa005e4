#!/usr/bin/env python3
"""bench.py — throughput of the Leon compression hot path (leon -c: count -> Bloom -> anchors -> walk -> range coding,
plus lossy quality smoothing and the quality blocks' zlib streams) on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (torchrun for N > 1)
    python bench.py --impl reference --steps K --warmup W     # the reference's own CPU implementation (oracle/_ref/leon)

Workload = BASELINE.json configs[2] (the configuration the metric is quoted on at 1/2/4/8 GPUs): 4.6 Mbp genome, 4.6 M x
100 bp FASTQ reads, lossy qualities, k = 31.  A step = one pass of the whole hot path over that read set.  `value` =
reads/s with the read set already resident in HBM; `e2e` = the same through the C ABI with pinned HOST buffers (H2D of
the reads and D2H of every output inside the timed region).  Both arms use the SAME full-size seeded read set; every
line carries `output_digest` / `digest_ok`: the digest of all output streams against the one committed from a run of the
unmodified reference with -nb-cores 1 on that exact read set (tests/golden/digest_*.json, tools/make_digest.py).
configs[1] (FASTA, seq-only) is measured as the extra key `configs1`.  One JSON line on stdout.
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from leon_b200 import workloads as W  # noqa: E402  (numpy only)

METRIC = "reads/s compressed (leon -c: count + bloom + anchors + walk + range coding + lossy quality stream)"
KMER = W.KMER
RPB = W.RPB
REF_WALL_BUDGET_S = 170.0      # reference arm: timed repetitions stop when this much wall time has been spent in them


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="configs[2]", choices=sorted(W.CONFIGS))
    ap.add_argument("--reads", type=int, default=0, help="override the number of reads (debug; marks the line invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the configs[1] extra measurement")
    ap.add_argument("--host-zlib", action="store_true", help="return the quality text instead of zlib streams written by the GPU")
    ap.add_argument("--anchor-batch", type=int, default=0, help="reads per anchor-resolution batch (0 = library default)")
    ap.add_argument("--phases", action="store_true", help="print per-phase / per-kernel device times to stderr")
    ap.add_argument("--no-large", action="store_true", help="skip the configs[3] streaming measurement (extra key `large`)")
    ap.add_argument("--large-reads", type=int, default=0, help="debug: reads of the configs[3]-shaped set (genome scaled with it)")
    ap.add_argument("--only-large", action="store_true", help="only the streaming measurement (configs[3], or --large-shape configs4)")
    ap.add_argument("--large-shape", default="configs3", choices=["configs3", "configs4"],
                    help="configs4: a 1/20 sample of BASELINE configs[4] (k = 63, lossless, 30 M x 150 bp from a 150 Mbp genome)")
    ap.add_argument("--large-genome", type=int, default=0, help="genome length of the streaming measurement (default: scaled with the reads)")
    ap.add_argument("--large-batch-chunks", type=int, default=32, help="chunks of 250 000 reads per rank and streaming batch")
    ap.add_argument("--large-passes", type=int, default=0, help="counting passes of the streaming measurement (0 = from free memory)")
    return ap.parse_args()


def config_block(name, reads):
    """The `config` object: identical in both arms (same workload, same size)."""
    c = W.CONFIGS[name]
    return {"workload": c["text"] if reads == c["reads"] else "DEBUG SIZE (invalid as a bench line): %d reads of %s" % (reads, name),
            "reads": reads, "read_len": c["read_len"], "kmer_size": KMER, "reads_per_block": RPB,
            "quality": "lossy" if c["fastq"] else "none (seq-only)",
            "l2": "inputs larger than L2 (460 MB of bases, 2.6 GB of keys per step); no flush needed"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def golden_digest(name):
    p = os.path.join(ROOT, "tests", "golden", "digest_%s.json" % name.replace("[", "").replace("]", ""))
    return json.load(open(p)) if os.path.exists(p) else None


# ---------------------------------------------------------------------------------------------------------------
# clocks sampling during the timed region
# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the UNMODIFIED reference binary (oracle/_ref/leon, built by oracle/build_ref.sh from
# /root/reference) on the host cores, on the SAME full-size read set.  Test infrastructure executed as the thing to
# compare with, never as product.
# ---------------------------------------------------------------------------------------------------------------
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "leon")


def reference_available():
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def write_workload_file(name, tmpdir, reads=0):
    c = W.CONFIGS[name]
    bases, offs, quals = W.generate(name, reads=reads)
    path = os.path.join(tmpdir, "w.fastq" if c["fastq"] else "w.fasta")
    W.write_fixed(path, bases, c["read_len"], quals)
    return path, len(offs) - 1


def reference_command(name, path, cores):
    return [REF_BIN, "-file", path, "-c", "-kmer-size", str(KMER), "-nb-cores", str(cores), "-verbose", "0"] + W.CONFIGS[name]["ref_flags"]


def run_reference_once(name, path, cores, workdir):
    """One `leon -c` run of the reference on the workload file; returns wall seconds."""
    for f in os.listdir(workdir):
        if f.endswith(".leon") or f.endswith(".h5") or f.startswith("trashme"):
            p = os.path.join(workdir, f)
            shutil.rmtree(p, ignore_errors=True) if os.path.isdir(p) else os.remove(p)
    t0 = time.perf_counter()
    r = subprocess.run(reference_command(name, path, cores), cwd=workdir, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    dt = time.perf_counter() - t0
    if r.returncode != 0 or not os.path.exists(path + ".leon"):
        raise RuntimeError("reference leon failed (rc %d): %s" % (r.returncode, r.stdout[-400:]))
    return dt


def cpu_baseline_block(name, reads):
    """The reference binary with all host cores on the full workload file: one timed run.  On the same (page-cached) file the
    drop-in itself: the `leon` command line tool of this repo as a process (CUDA context creation, file parse, GPU path,
    container write: everything a user of `leon -file X -c` waits for) and the same call inside this process
    (leon_compress_file with the context already up), so that the ratio of the actual drop-in is visible (`file_e2e`)."""
    cores = os.cpu_count() or 1
    if not reference_available():
        return {"value": None, "unit": "reads/s", "cores": cores, "kind": "reference", "sample": "oracle/_ref/leon missing on this box"}, None
    tmp = tempfile.mkdtemp(prefix="leonref_")
    try:
        path, n = write_workload_file(name, tmp, reads)
        t = run_reference_once(name, path, cores, tmp)
        cpu = {"value": n / t, "unit": "reads/s", "cores": cores, "kind": "reference", "same_config": True,
               "sample": "the full workload (%d reads), one run of `%s`, wall %.2f s incl. file read" % (
                   n, " ".join(["leon"] + reference_command(name, "X", cores)[3:]), t)}
        file_e2e = None
        try:
            import leon_b200
            cli = os.path.join(ROOT, "leon_b200", "bin", "leon")
            flags = W.CONFIGS[name]["ref_flags"]
            out = path + ".b200.leon"
            walls = []
            for _ in range(2):
                t0 = time.perf_counter()
                env = {k: v for k, v in os.environ.items() if k != "CUDA_DEVICE_MAX_CONNECTIONS"}     # what a user's shell has
                r = subprocess.run([cli, "-file", path, "-c", "-kmer-size", str(KMER), "-verbose", "0", "-out", out] + flags,
                                   stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
                walls.append(time.perf_counter() - t0)
                if r.returncode != 0:
                    raise RuntimeError(r.stdout[-300:])
            inproc = []
            for _ in range(2):
                t0 = time.perf_counter()
                leon_b200.compress_file(path, out, kmer_size=KMER, noheader="-noheader" in flags, seq_only="-seq-only" in flags)
                inproc.append(time.perf_counter() - t0)
            file_e2e = {"what": "wall clock of the whole drop-in on the page-cached workload file (names not kept, like the reference arm): "
                                "process = `leon -file X -c` of this repo incl. CUDA context creation; in_process = leon_compress_file "
                                "with the context already up; reference = oracle/_ref/leon with all cores, same file",
                        "process_s": min(walls), "in_process_s": min(inproc), "reference_s": t,
                        "process_reads_per_s": n / min(walls), "in_process_reads_per_s": n / min(inproc),
                        "ratio_process": t / min(walls), "ratio_in_process": t / min(inproc),
                        "output_bytes": os.path.getsize(out), "reference_output_bytes": os.path.getsize(path + ".leon")}
        except Exception as e:       # the figure is an extra; never lose the line over it
            file_e2e = {"error": str(e)[:300]}
        return cpu, file_e2e
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not reference_available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/leon not built on this box"}))
        return 0
    name = args.workload
    cores = os.cpu_count() or 1
    tmp = tempfile.mkdtemp(prefix="leonref_")
    try:
        path, n = write_workload_file(name, tmp, args.reads)
        warm = []
        for _ in range(min(args.warmup, 1)):           # one warm-up run brings the file into the page cache; more would only
            warm.append(run_reference_once(name, path, cores, tmp))   # spend the wall budget (every run is ~10 s of 16 cores)
        t = []
        while len(t) < args.steps and (not t or sum(t) + t[-1] <= REF_WALL_BUDGET_S):
            t.append(run_reference_once(name, path, cores, tmp))
        digest = W.digest_of_container(path + ".leon", W.CONFIGS[name]["fastq"])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    total = sum(t)
    value = n * len(t) / total
    gold = golden_digest(name)
    # with more than one thread the reference's anchor order (and with it dna / anchorsDict) is not deterministic; the
    # Bloom filter and the quality text are
    same = {k: (gold is not None and digest.get(k) == gold.get(k)) for k in ("bloom", "qual_text") if k in digest}
    sample = "the full workload, %d timed runs of %d requested (wall budget %.0f s), %d warm-up" % (len(t), args.steps, REF_WALL_BUDGET_S, len(warm))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "steps_timed": len(t), "warmup_run": len(warm), "ms_per_step": 1000.0 * total / len(t),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": config_block(name, n),
        "command": " ".join(["leon"] + reference_command(name, "X", cores)[3:]),
        "input_gbs": value * W.CONFIGS[name]["read_len"] / 1e9,
        "cpu_baseline": {"value": value, "unit": "reads/s", "cores": cores, "kind": "reference", "sample": sample, "same_config": True},
        "e2e": {"value": value, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "deterministic_streams_equal_nb_cores_1": same,
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------------------
# this repo's arm
# ---------------------------------------------------------------------------------------------------------------
def algorithmic_bytes(kernel, st):
    """Algorithmic bytes ONE launch of `kernel` moves (DESIGN.md 'kernels and rooflines').
    st: nreads, nbases, nkmers, nruns, nsym, launches (per step, of this kernel)."""
    R, NB, NK = st["nreads"], st["nbases"], st["nkmers"]
    kb = 8 if KMER <= 31 else 16
    n = max(1.0, st.get("launches", {}).get(kernel, 1.0))      # kernels launched once per anchor batch share the totals
    if kernel == "k_walk":            # one 544-byte Bloom window per extension step: (L - k) steps per read
        return 544.0 * (NB - KMER * R) / n
    if kernel == "k_msd_scatter":     # read every key once, write it once (one launch per 8-bit level of the hash partition)
        return 2.0 * NK * kb
    if kernel == "k_msd_hist":
        return 1.0 * NK * kb
    if kernel == "k_part_scatter":    # stable radix pass over the 8-byte symbol records
        return 2.0 * st["nsym"] * 8.0 / n * 2.0
    if kernel == "k_part_hist":
        return 1.0 * st["nsym"] * 8.0 / n * 2.0
    if kernel == "k_bucket_count":    # read every key once, write (key, count) of the solid ones; averaged over its two launches
        return (1.0 * NK * kb + st["nruns"] * (kb + 4.0)) / 2.0
    if kernel == "k_extract":         # read the bases, write one key per k-mer
        return NB + NK * kb
    if kernel == "k_chunk_ranks":     # read the 8-byte sorted record, write the 17-byte coder record; histogram row per chunk
        return st["nsym"] * (8.0 + 17.0 + 2.0) / n
    if kernel == "k_emit_syms":
        return (st["nsym"] * 8.0 + NB) / n
    if kernel == "k_range_encode":    # read the 17-byte record of every symbol, write the output bytes
        return (st["nsym"] * 17.0 + st["dna_bytes"]) / n
    if kernel == "k_deflate_count":   # read the quality text once (frequencies and Adler sums are small)
        return st["qual_bytes"]
    if kernel == "k_deflate_emit":    # read the quality text once, write the compressed chunk once
        return st["qual_bytes"] + st.get("qual_z_bytes", 0.0)
    if kernel == "k_deflate_gather":  # move the compressed chunks to their place
        return 2.0 * st.get("qual_z_bytes", 0.0)
    if kernel == "k_quals":           # read bases + qualities, write the smoothed text (Bloom lookups not counted)
        return 2.0 * NB + st["qual_bytes"]
    return None


def measure(args, name, torch, dist, world, rank, local_rank, want_e2e=True):
    """Resident and end-to-end measurement of one workload; returns a dict (rank 0: complete; other ranks: partial)."""
    import leon_b200
    from leon_b200 import shard
    c = W.CONFIGS[name]
    fastq = c["fastq"]
    read_len = c["read_len"]
    total_reads = args.reads if args.reads else c["reads"]

    # ---- the workload: every rank generates the same seeded read set; with N > 1 ranks the set is dealt in rounds of
    # N * S consecutive reads, S to each rank in rank order (leon_b200.shard; include/leon_b200.h leon_comm_*), so the
    # total work is fixed as N grows ("strong").
    S = 0
    if world == 1:
        pin_b = leon_b200.api.PinnedArray(total_reads * read_len)
        pin_q = leon_b200.api.PinnedArray(total_reads * read_len) if fastq else None
        bases, offs, quals = W.generate(name, out_bases=pin_b.array, out_quals=pin_q.array if fastq else None, reads=total_reads)
    else:
        S = args.anchor_batch if args.anchor_batch else shard.choose_batch_reads(total_reads, world, RPB)
        fb, fo, fq = W.generate(name, reads=total_reads)
        sb, offs, sq = shard.shard_reads(fb, fo, fq, rank, world, S)
        del fb, fq
        pin_b = leon_b200.api.PinnedArray(max(1, len(sb)))
        pin_b.array[:len(sb)] = sb
        bases = pin_b.array[:len(sb)]
        quals = None
        if fastq:
            pin_q = leon_b200.api.PinnedArray(max(1, len(sq)))
            pin_q.array[:len(sq)] = sq
            quals = pin_q.array[:len(sq)]
        del sb, sq
    local_reads = len(offs) - 1
    pin_o = leon_b200.api.PinnedArray((local_reads + 1) * 8, np.uint64)
    pin_o.array[:] = offs
    offs = pin_o.array

    dev_zlib = fastq and not args.host_zlib
    eng = leon_b200.Leon(kmer_size=KMER, reads_per_block=RPB, has_qual=fastq, device=local_rank,
                         anchor_batch=S if world > 1 else args.anchor_batch, qual_zlib_on_device=dev_zlib, skip_qual_text=dev_zlib,
                         skip_read_arrays=True)
    stream = torch.cuda.current_stream()
    eng.set_stream(stream.cuda_stream)
    if world > 1:
        uid = [leon_b200.Leon.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        eng.comm_init(rank, world, uid[0])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: inputs resident in HBM
    eng.upload(bases, offs, quals)
    for _ in range(args.warmup):
        eng.run_device()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = eng.timings()["kernel_launches"]
    e0.record(stream)
    ktimes, phase = {}, {}
    for _ in range(args.steps):
        eng.run_device()
        for k, (ms, n) in eng.kernel_times().items():
            a = ktimes.setdefault(k, [0.0, 0])
            a[0] += ms
            a[1] += n
        for k, v in eng.timings().items():
            if k.endswith("_ms") or k in ("anchor_batches", "anchor_rounds"):
                phase[k] = phase.get(k, 0.0) + v
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    launches = eng.timings()["kernel_launches"] - launches0
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([dev_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())
    ms_per_step = dev_ms / args.steps
    value = total_reads / (ms_per_step / 1000.0)
    res = eng.fetch(copy=True)

    # ---- output digest: all ranks' outputs merged in file order on rank 0, against the committed reference digest
    digest, digest_ok, digest_diff = None, None, None
    allr = [res]
    if world > 1:
        allr = [None] * world if rank == 0 else None
        dist.gather_object(res, allr, dst=0)
    if rank == 0:
        merged = res if world == 1 else shard.merge_results(allr, total_reads, world, S, RPB)
        digest = W.digest_of_result(merged, fastq)
        gold = golden_digest(name)
        if gold is not None and total_reads == gold.get("reads"):
            digest_ok, digest_diff = W.digests_equal(digest, gold)
    del allr

    # the roofline describes rank 0's launches: its own share of the reads
    stats = {"nreads": local_reads, "nbases": local_reads * read_len, "nkmers": local_reads * (read_len - KMER + 1),
             "nruns": res["nb_distinct"], "dna_bytes": len(res["dna"]),
             "qual_bytes": float(local_reads * (read_len + 1)) if fastq else 0.0, "qual_z_bytes": float(len(res["qual_z"])),
             "launches": {k: v[1] / args.steps for k, v in ktimes.items()}}
    stats["nsym"] = float(eng.timings()["nsym_total"])
    out = dict(name=name, total_reads=total_reads, local_reads=local_reads, read_len=read_len, S=S, value=value, ms_per_step=ms_per_step,
               dev_ms=dev_ms, launches=launches, clocks=clocks, ktimes=ktimes, phase=phase, stats=stats, digest=digest,
               digest_ok=digest_ok, digest_diff=digest_diff, anchors=res["anchor_count"], dna_bytes=len(res["dna"]),
               bloom_bytes=int(len(res["bloom"])), cutoff=res["cutoff"], solid=res["nb_solids_cutoff"])

    # ---- e2e: host buffers in, host buffers out, through the C ABI one-shot call
    if want_e2e:
        for _ in range(min(args.warmup, 2)):
            eng.compress(bases, offs, quals, copy=False)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            r = eng.compress(bases, offs, quals, copy=False)
        g1.record(stream)
        barrier()
        e2e_wall = (time.perf_counter() - t0) * 1000.0
        e2e_ms = max(g0.elapsed_time(g1), e2e_wall) / args.steps
        if world > 1:
            t = torch.tensor([e2e_ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms = float(t.item())
        h2d = int(bases.nbytes + offs.nbytes + (quals.nbytes if quals is not None else 0))
        d2h = int(len(r["bloom"]) + len(r["anchors_dict"]) + len(r["dna"]) + r["blocksizes"].nbytes +
                  r["anchor_pos"].nbytes + r["anchor_addr"].nbytes + len(r["qual_text"]) + len(r["qual_z"]))
        if world > 1:   # bytes of all ranks
            t = torch.tensor([h2d, d2h, launches], device="cuda", dtype=torch.int64)
            dist.all_reduce(t)
            h2d, d2h, out["launches"] = (int(x) for x in t.tolist())
        out.update(e2e_ms=e2e_ms, e2e_value=total_reads / (e2e_ms / 1000.0), h2d=h2d, d2h=d2h)
    eng.close()
    pin_b.free()
    pin_o.free()
    if fastq:
        pin_q.free()
    return out


def measure_large(args, torch, dist, world, rank, local_rank):
    """BASELINE configs[3] (100 Mbp genome, 33.3 M x 150 bp FASTQ with N runs and indels, Bloom filter 150 MB > L2) through the
    STREAMING entry points from pinned host memory on `world` ranks: every rank is dealt the chunks (250 000 reads) c with
    c % world == rank, feeds them in batches, and hashes its own blocks; rank 0 combines the per-block digests."""
    import concurrent.futures as cf
    import hashlib
    import zlib
    import leon_b200
    from leon_b200 import synth
    name = "configs[3]"
    c = dict(W.CONFIGS[name])
    kmer, lossless = KMER, False
    if args.large_shape == "configs4":
        # BASELINE configs[4] (3 Gbp, 600 M x 150 bp FASTQ at 30x, k = 63, lossless qualities, streamed) at 1/20 of its size by
        # default: same read shape and coverage, genome and read count scaled together
        name = "configs[4]"
        frac = 20
        c.update(genome=3_000_000_000 // frac, reads=600_000_000 // frac, gseed=5, rseed=6,
                 text="configs[4] at 1/%d scale: %d Mbp random genome, 150 bp FASTQ reads at 30x, N bases and N runs, indels, k=63, "
                      "lossless qualities, streamed in batches" % (frac, 3000 // frac))
        kmer, lossless = 63, True
    total_reads = args.large_reads if args.large_reads else c["reads"]
    genome_len = c["genome"] if not args.large_reads else max(1_000_000, int(c["genome"] * total_reads / c["reads"]))
    if args.large_genome:
        genome_len = args.large_genome
    S = W.C3_CHUNK
    nchunks = (total_reads + S - 1) // S
    mine = [ch for ch in range(nchunks) if ch % world == rank]
    t0 = time.perf_counter()
    g = synth.make_genome(genome_len, c["gseed"])
    threads = max(1, min(16, (os.cpu_count() or 1) // world))
    with cf.ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(lambda ch: W.c3_chunk(g, ch, total_reads, c["read_len"], c["rseed"], True), mine))
    lens = np.concatenate([p[1] for p in parts]) if parts else np.zeros(0, np.int64)
    nb = int(lens.sum())
    pin_b, pin_q = leon_b200.api.PinnedArray(max(1, nb)), leon_b200.api.PinnedArray(max(1, nb))
    pin_o = leon_b200.api.PinnedArray((len(lens) + 1) * 8, np.uint64)
    at = 0
    for p in parts:
        pin_b.array[at:at + len(p[0])] = p[0]
        pin_q.array[at:at + len(p[0])] = p[2]
        at += len(p[0])
    pin_o.array[0] = 0
    np.cumsum(lens, out=pin_o.array[1:])
    del parts, g
    bases, quals, offs = pin_b.array[:nb], pin_q.array[:nb], pin_o.array
    gen_s = time.perf_counter() - t0
    local_reads = len(lens)
    # batches of M chunks per rank; every rank makes the same number of calls (empty batches at the end where it has run out)
    from leon_b200 import shard
    M = max(1, args.large_batch_chunks)
    batches = shard.stream_batches(total_reads, rank, world, S, M)
    ncalls = len(batches)
    assert local_reads == sum(hi - lo for lo, hi in shard.rank_read_ranges(total_reads, rank, world, S))

    torch.cuda.synchronize()
    free_start = torch.cuda.mem_get_info()[0]
    eng = leon_b200.Leon(kmer_size=kmer, reads_per_block=RPB, has_qual=True, lossless=lossless, device=local_rank, anchor_batch=S,
                         qual_zlib_on_device=True, skip_qual_text=True, skip_read_arrays=True)

    class MemSampler(threading.Thread):      # peak device memory of this rank while the runs are going
        def __init__(self):
            super().__init__(daemon=True)
            self.min_free, self.stop = free_start, False

        def run(self):
            while not self.stop:
                self.min_free = min(self.min_free, torch.cuda.mem_get_info()[0])
                time.sleep(0.01)
    mem = MemSampler()
    mem.start()
    stream = torch.cuda.current_stream()
    eng.set_stream(stream.cuda_stream)
    if world > 1:
        uid = [leon_b200.Leon.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        eng.comm_init(rank, world, uid[0])

    cuts = []
    for lo, hi in batches:              # the arrays of every batch once: a prefetched batch is recognised by its pointers
        b0, b1 = int(offs[lo]), int(offs[hi])
        cuts.append((bases[b0:b1], np.ascontiguousarray(offs[lo:hi + 1] - offs[lo]), quals[b0:b1]))

    def one_run(keep):
        phases = {}
        t = time.perf_counter()
        npass = eng.count_begin(local_reads, nb, args.large_passes)
        for ps in range(npass):
            for i, (b, o, _q) in enumerate(cuts):
                if i + 1 < len(cuts):
                    eng.prefetch(cuts[i + 1][0], cuts[i + 1][1])
                elif ps + 1 < npass:
                    eng.prefetch(cuts[0][0], cuts[0][1])        # first batch of the next counting pass
                else:
                    eng.prefetch(*cuts[0])                      # first batch of the encoding pass (with its qualities)
                eng.count_batch(b, o, i == len(cuts) - 1)
        cnt = eng.count_finish()
        phases["count_s"] = time.perf_counter() - t
        if keep:        # CUDA-event time per kernel of the counting pass(es), ms (the timers restart at count_begin)
            phases["count_kernels_ms"] = {k: round(v[0], 2) for k, v in sorted(eng.kernel_times().items(), key=lambda kv: -kv[1][0])[:8]}
        t = time.perf_counter()
        eng.bloom_build()
        phases["bloom_s"] = time.perf_counter() - t
        t = time.perf_counter()
        first = 0
        outs = []
        per_batch = []
        for i, (b, o, q) in enumerate(cuts):
            if i + 1 < len(cuts):
                eng.prefetch(*cuts[i + 1])
            tb = time.perf_counter()
            bo = eng.encode_block(first, b, o, q, copy=keep)
            first += bo["nblocks"]
            if keep:
                outs.append(bo)
                tmb = eng.timings()     # CUDA-event phases of this batch beside the wall clock of the call
                per_batch.append({"call_ms": round((time.perf_counter() - tb) * 1e3, 1), "anchor_ms": round(tmb["anchor_ms"], 1),
                                  "walk_entropy_ms": round(tmb["walk_ms"], 1), "reads": int(len(o) - 1)})
        res = eng.encode_finish(copy=keep)
        phases["encode_s"] = time.perf_counter() - t
        if keep:        # phases of the LAST batch (CUDA events), main-stream kernels of the whole run + walk / entropy kernels of the last batch
            phases["batches_ms"] = per_batch
            phases["kernels_ms_cumulative_and_last_batch"] = {k: round(v[0], 2) for k, v in sorted(eng.kernel_times().items(), key=lambda kv: -kv[1][0])[:14]}
        tm = eng.timings()
        # the anchor dictionary is ONE never-reset range coder (C/Leon.cpp:1245-1260): a serial chain on a host thread of rank 0
        # that runs beside the whole encoding pass; dict_wait = how long the pass then still waited for it
        phases["dict_host_s"] = tm["dict_host_ms"] / 1e3
        phases["dict_wait_s"] = tm["dict_wait_ms"] / 1e3
        return npass, cnt, outs, res, phases

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    one_run(False)                       # warm-up (allocations)
    steps = max(1, min(args.steps, 2))
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    w0 = time.perf_counter()
    for i in range(steps):
        npass, cnt, outs, res, phases = one_run(i == steps - 1)
    e1.record(stream)
    barrier()
    ms = max(e0.elapsed_time(e1), (time.perf_counter() - w0) * 1000.0) / steps
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # ---- digest: per-block checksums on every rank, combined in file order on rank 0
    def block_hashes(bo):
        out = []
        da, za = 0, 0
        for j in range(bo["nblocks"]):
            dn, zn = int(bo["blocksizes"][2 * j]), int(bo["qual_z_block_sizes"][j])
            out.append((hashlib.sha256(bo["dna"][da:da + dn]).digest(), hashlib.sha256(zlib.decompress(bo["qual_z"][za:za + zn])).digest(),
                        int(bo["blocksizes"][2 * j]), int(bo["blocksizes"][2 * j + 1])))
            da += dn
            za += zn
        return out
    with cf.ThreadPoolExecutor(threads) as ex:
        per = [h for lst in ex.map(block_hashes, outs) for h in lst]
    # local block j of chunk-round i on this rank is global block (i * world + rank) * (S / RPB) + j
    gids = shard.global_block_ids(total_reads, rank, world, S, RPB)
    assert len(gids) == len(per), (len(gids), len(per))
    mem.stop = True
    mem.join()
    peak_bytes = int(free_start - mem.min_free)
    if world > 1:
        t = torch.tensor([peak_bytes], device="cuda", dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        peak_bytes = int(t.item())
    mine_out = dict(gids=gids, per=per, bloom=hashlib.sha256(res["bloom"]).hexdigest() if rank == 0 else None,
                    dict=hashlib.sha256(res["anchors_dict"]).hexdigest() if rank == 0 else None, anchors=res["anchor_count"])
    allo = [mine_out]
    if world > 1:
        allo = [None] * world if rank == 0 else None
        dist.gather_object(mine_out, allo, dst=0)
    eng.close()
    pin_b.free(); pin_q.free(); pin_o.free()
    if rank != 0:
        return None
    table = {}
    for o in allo:
        for gid, h in zip(o["gids"], o["per"]):
            table[gid] = h
    order = sorted(table)
    bs = np.array([[table[b][2], table[b][3]] for b in order], np.uint64).reshape(-1)
    digest = {"bloom": allo[0]["bloom"], "anchors_dict": allo[0]["dict"], "blocksizes": W._sha(bs), "nblocks": len(order),
              "dna_nbytes": int(bs[0::2].sum()),
              "dna_cc": hashlib.sha256(b"".join(table[b][0] for b in order)).hexdigest(),
              "qual_text_cc": hashlib.sha256(b"".join(table[b][1] for b in order)).hexdigest()}
    gold = golden_digest(name) if (total_reads == c["reads"] and name == "configs[3]") else None
    ok, diff = None, None
    if gold is not None and "dna_cc" in gold:
        keys = ["bloom", "anchors_dict", "blocksizes", "dna_cc", "qual_text_cc"]
        diff = [k for k in keys if digest.get(k) != gold.get(k)]
        ok = not diff
    value = total_reads / (ms / 1000.0)
    return {"workload": c["text"] if total_reads == c["reads"] else "DEBUG SIZE: %d reads of configs[3]" % total_reads,
            "api": "streaming entry points (leon_count_begin / leon_count / leon_count_finish / leon_bloom_build / leon_encode_block / "
                   "leon_encode_finish), host buffers in, host buffers out: this is an end-to-end figure",
            "value": value, "unit": "reads/s", "ms_per_step": ms, "steps": steps, "n_gpus": world, "count_passes": npass,
            "kmer_size": kmer, "quality": "lossless" if lossless else "lossy", "reads": total_reads, "genome": genome_len,
            "peak_device_bytes_per_rank": peak_bytes, "host_bytes_resident_per_rank": int(2 * nb),
            "batches_per_rank": ncalls, "reads_per_batch_per_rank": M * S,
            "input_gbs": value * c["read_len"] * 2 / 1e9, "phases_s_rank0": phases, "anchors": allo[0]["anchors"],
            "cutoff": cnt["cutoff"], "solid": cnt["nb_solids_cutoff"], "bloom_bytes": int(len(res["bloom"])),
            "output_digest": digest, "digest_ok": ok, "digest_mismatch": diff,
            "digest_reference": "tests/golden/digest_configs3.json: oracle/_ref/leon -nb-cores 1 on the same seeded read set "
                                "(dna_cc / qual_text_cc = checksum of the per-block checksums)",
            "generation_s": gen_s}


def main_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torchrun (one rank per GPU)" % args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    name = args.workload
    if args.only_large:
        lg = measure_large(args, torch, dist, world, rank, local_rank)
        if rank == 0:
            print(json.dumps({"large": lg}))
        return 0
    m = measure(args, name, torch, dist, world, rank, local_rank)
    extra = None
    if not args.no_extra and name != "configs[1]" and not args.reads:
        extra = measure(args, "configs[1]", torch, dist, world, rank, local_rank, want_e2e=False)

    ktimes, stats, clocks = m["ktimes"], m["stats"], m["clocks"]
    read_len = m["read_len"]
    # ---- roofline of the dominant kernel (largest share of the device time of the timed region)
    peak, peak_src = peaks()
    dom = max(ktimes.items(), key=lambda kv: kv[1][0])[0] if ktimes else None
    roofline = None
    if dom:
        ms_tot, n = ktimes[dom]
        per_launch_ms = ms_tot / max(n, 1)
        ab = algorithmic_bytes(dom, stats)
        ach = (ab / 1e9) / (per_launch_ms / 1e3) if ab else None
        roofline = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                    "frac": (ach / peak) if ach else None, "traffic": None, "peak_source": peak_src,
                    "launch_ms": per_launch_ms, "launches_per_step": n / args.steps,
                    "share_of_step": ms_tot / m["dev_ms"], "algorithmic_bytes_per_launch": ab}
        if dom == "k_range_encode":
            # one warp per block stream runs the serial low/range recurrence: bound by single-warp dependent-issue latency,
            # not by bytes -- the figure that describes it is cycles per symbol of a stream
            nblk = max(1, -(-m["local_reads"] // RPB))
            sym_per_block = stats["nsym"] / nblk
            roofline["note"] = "latency-bound serial recurrence (one warp per 50000-read block); HBM fraction is not its limit; " \
                               "launches of different anchor batches overlap each other and the other kernels"
            roofline["ns_per_symbol"] = per_launch_ms * 1e6 / (sym_per_block)
            if clocks.get("sm_mhz"):
                roofline["cycles_per_symbol"] = roofline["ns_per_symbol"] * clocks["sm_mhz"] / 1e3
        tr = os.path.join(ROOT, "profiles", "traffic.json")   # dram bytes per launch from the committed ncu --set full capture
        if os.path.exists(tr):
            try:
                roofline["traffic"] = json.load(open(tr)).get(dom)
            except Exception:
                pass
    kernel_rooflines = {}
    for kname, (ms_tot_k, n_k) in ktimes.items():
        ab_k = algorithmic_bytes(kname, stats)
        if ab_k and n_k and kname != "k_walk":     # k_walk: see probe_roofline (its filter is L2-resident here; an HBM fraction is meaningless)
            gbs = (ab_k / 1e9) / ((ms_tot_k / n_k) / 1e3)
            kernel_rooflines[kname] = {"ms_per_launch": ms_tot_k / n_k, "launches_per_step": n_k / args.steps,
                                       "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / peak}
    # Bloom-probe roofline of k_walk (north_star): extension steps (= contains4 windows) per second against the plateau of
    # an independent microbenchmark with the kernel's access shape (tools/probe_roofline.cu, measured on B200)
    probe = None
    if "k_walk" in ktimes and ktimes["k_walk"][1]:
        wms = ktimes["k_walk"][0] / args.steps
        steps = float(stats["nbases"] - KMER * stats["nreads"])
        wps = steps / (wms / 1e3)
        probe = {"kernel": "k_walk", "windows_per_s": wps, "filter_bytes": m["bloom_bytes"], "peak_windows_per_s": None,
                 "frac": None, "regime": None}
        try:
            prj = json.load(open(os.path.join(ROOT, "profiles", "probe_roofline.json")))
            pr = prj["windows_per_s"]
            fb = m["bloom_bytes"]
            key = "l2_resident_6.9MB" if fb <= 64e6 else ("150MB" if fb <= 300e6 else ("dram_1GB" if fb <= 2e9 else "dram_4.5GB"))
            probe.update(peak_windows_per_s=pr[key], frac=wps / pr[key], regime=key, peak_source=prj.get("source", "tools/probe_roofline.cu microbenchmark"),
                         ncu_context=prj.get("k_walk_ncu"))
        except Exception:
            pass
    if args.phases and rank == 0:
        print("phases (ms/step):", {k: round(v / args.steps, 3) for k, v in m["phase"].items()}, file=sys.stderr)
        print("kernels (ms/step, launches/step):",
              {k: (round(v[0] / args.steps, 3), v[1] // args.steps) for k, v in sorted(ktimes.items(), key=lambda kv: -kv[1][0])},
              file=sys.stderr)
        print("anchors %d, dna bytes %d, symbols %d, solid %d, cutoff %d" % (
            m["anchors"], m["dna_bytes"], int(stats["nsym"]), m["solid"], m["cutoff"]), file=sys.stderr)

    large = None
    if not args.no_large and not args.reads:
        large = measure_large(args, torch, dist, world, rank, local_rank)
    cpu, file_e2e = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu, file_e2e = cpu_baseline_block(name, args.reads)
    if rank != 0:
        return 0
    line = {
        "metric": METRIC, "value": m["value"], "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic",
        "config": config_block(name, m["total_reads"]),
        "parallelism": "1 process per GPU" if world == 1 else
        "%d ranks, reads dealt in rounds of %d x %d; NCCL: k-mer all-to-all by hash owner, Bloom OR as reduce-scatter + "
        "all-gather, anchor candidates all-gathered per batch" % (world, world, m["S"]),
        "input_gbs": m["value"] * read_len * (2 if W.CONFIGS[name]["fastq"] else 1) / 1e9,
        "e2e": {"value": m["e2e_value"], "unit": "reads/s", "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"],
                "ms_per_step": m["e2e_ms"], "input_gbs": m["e2e_value"] * read_len * (2 if W.CONFIGS[name]["fastq"] else 1) / 1e9},
        "gpu_launches": int(m["launches"]),
        "clocks": clocks,
        "output_digest": m["digest"], "digest_ok": m["digest_ok"], "digest_mismatch": m["digest_diff"],
        "digest_reference": "tests/golden/digest_%s.json: oracle/_ref/leon -nb-cores 1 on the same seeded read set" % name.replace("[", "").replace("]", ""),
        "roofline": roofline,
        "kernel_rooflines": kernel_rooflines,
        "probe_roofline": probe,
        "cpu_baseline": cpu,
        "file_e2e": file_e2e,
        "phases_ms_per_step": {k: v / args.steps for k, v in m["phase"].items()},
        "kernels_ms_per_step": {k: v[0] / args.steps for k, v in ktimes.items()},
    }
    if extra is not None:
        line["configs1"] = {"workload": W.CONFIGS["configs[1]"]["text"], "value": extra["value"], "unit": "reads/s",
                            "ms_per_step": extra["ms_per_step"], "digest_ok": extra["digest_ok"], "digest_mismatch": extra["digest_diff"],
                            "phases_ms_per_step": {k: v / args.steps for k, v in extra["phase"].items()}}
    if large is not None:
        line["large"] = large
    print(json.dumps(line))
    return 0


if __name__ == "__main__":
    a = parse_args()
    rc = main_reference(a) if a.impl == "reference" else main_b200(a)
    if a.impl != "reference" and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as _dist
        if _dist.is_initialized():
            _dist.barrier()
            _dist.destroy_process_group()
    sys.exit(rc)
