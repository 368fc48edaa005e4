#!/usr/bin/env python3
"""bench.py — throughput of the Leon compression hot path (leon -c sequence stream) on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W     # the reference's own CPU implementation (oracle/_ref/leon)

A step = one pass of the whole hot path (k-mer counting -> Bloom build -> anchors -> walk -> range coding) over the
BASELINE.json configs[1] read set (4.6 Mbp random genome, 100 bp reads at 100x, 0.5 % substitutions, k=31,
sequence only).  `value` = reads/s with the read set already resident in HBM; `e2e` = the same through the C ABI with
pinned HOST buffers (H2D of the reads and D2H of every output inside the timed region).  One JSON line on stdout.
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

METRIC = "reads/s compressed (leon -c sequence stream: count + bloom + anchors + walk + range coding)"
GENOME_LEN = 4_600_000
READ_LEN = 100
COVERAGE = 100
KMER = 31
RPB = 50000
WORKLOAD = "configs[1]: 4.6 Mbp random genome, 100 bp FASTA reads at 100x, 0.5% substitutions, k=31, seq-only"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--reads", type=int, default=0, help="override the number of reads per GPU (debug; marks the line invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fastq", action="store_true", help="configs[2] shape: FASTQ with synthetic qualities, lossy (debug line)")
    ap.add_argument("--host-zlib", action="store_true", help="with --fastq: return the quality text instead of zlib streams written by the GPU")
    ap.add_argument("--anchor-batch", type=int, default=0, help="reads per anchor-resolution batch (0 = library default)")
    ap.add_argument("--phases", action="store_true", help="print per-phase / per-kernel device times to stderr")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


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
# /root/reference) on the host cores.  Test infrastructure executed as the thing to compare with, never as product.
# ---------------------------------------------------------------------------------------------------------------
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "leon")
SAMPLE_FRACTION = 0.1   # genome and read count scaled together: same 100x coverage, same read shape


def make_sample_fasta(tmpdir, fraction=SAMPLE_FRACTION):
    from leon_b200 import synth
    glen = int(GENOME_LEN * fraction)
    nreads = glen * COVERAGE // READ_LEN
    g = synth.make_genome(glen, 1)
    bases, offs, _ = synth.sample_reads(g, nreads, READ_LEN, 2)
    path = os.path.join(tmpdir, "sample.fasta")
    synth.write_fasta(path, bases, offs)
    return path, nreads, glen


def run_reference_once(fasta, cores, workdir):
    """One `leon -c -kmer-size 31 -seq-only -nb-cores <cores>` run; returns wall seconds."""
    for f in os.listdir(workdir):
        if f.endswith(".leon") or f.endswith(".h5") or f.startswith("trashme"):
            p = os.path.join(workdir, f)
            shutil.rmtree(p, ignore_errors=True) if os.path.isdir(p) else os.remove(p)
    t0 = time.perf_counter()
    r = subprocess.run([REF_BIN, "-file", fasta, "-c", "-kmer-size", str(KMER), "-seq-only", "-nb-cores", str(cores),
                        "-verbose", "0"], cwd=workdir, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    dt = time.perf_counter() - t0
    if r.returncode != 0 or not os.path.exists(fasta + ".leon"):
        raise RuntimeError("reference leon failed (rc %d): %s" % (r.returncode, r.stdout[-400:]))
    return dt


def reference_available():
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def cpu_baseline_block(runs=1):
    """Bounded sample of the same workload through the reference binary with all host cores."""
    cores = os.cpu_count() or 1
    if not reference_available():
        return {"value": None, "unit": "reads/s", "cores": cores, "kind": "reference",
                "sample": "oracle/_ref/leon missing on this box"}
    tmp = tempfile.mkdtemp(prefix="leonref_")
    try:
        fasta, nreads, glen = make_sample_fasta(tmp)
        best = min(run_reference_once(fasta, cores, tmp) for _ in range(runs))
        return {"value": nreads / best, "unit": "reads/s", "cores": cores, "kind": "reference",
                "sample": "%d reads x %d bp at %dx of a %d bp genome (1/%d of the workload, same coverage), "
                          "leon -c -kmer-size %d -seq-only -nb-cores %d, wall %.2f s incl. file read" % (
                              nreads, READ_LEN, COVERAGE, glen, round(1 / SAMPLE_FRACTION), KMER, cores, best)}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not reference_available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/leon not built on this box"}))
        return 0
    cores = os.cpu_count() or 1
    tmp = tempfile.mkdtemp(prefix="leonref_")
    try:
        fasta, nreads, glen = make_sample_fasta(tmp)
        for _ in range(args.warmup):
            run_reference_once(fasta, cores, tmp)
        t = [run_reference_once(fasta, cores, tmp) for _ in range(args.steps)]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    total = sum(t)
    value = nreads * args.steps / total
    sample = "each step = %d reads x %d bp at %dx of a %d bp genome (1/%d of the workload, same coverage)" % (
        nreads, READ_LEN, COVERAGE, glen, round(1 / SAMPLE_FRACTION))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * total / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "kmer_size": KMER, "reads_per_block": RPB, "sample": sample,
                   "command": "leon -c -kmer-size %d -seq-only -nb-cores %d" % (KMER, cores)},
        "input_gbs": value * READ_LEN / 1e9,
        "cpu_baseline": {"value": value, "unit": "reads/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------------------
# this repo's arm
# ---------------------------------------------------------------------------------------------------------------
def algorithmic_bytes(kernel, st):
    """Algorithmic bytes ONE launch of `kernel` moves (DESIGN.md 'kernels and rooflines').
    st: nreads, nbases, nkmers, nruns, nsym."""
    R, NB, NK = st["nreads"], st["nbases"], st["nkmers"]
    kb = 8 if KMER <= 31 else 16
    if kernel == "k_walk":            # one 544-byte Bloom window per extension step: (L - k) steps per read
        return 544.0 * (NB - KMER * R)
    if kernel == "k_msd_scatter":     # read every key once, write it once (one launch per 8-bit level of the hash partition)
        return 2.0 * NK * kb
    if kernel == "k_msd_hist":
        return 1.0 * NK * kb
    if kernel == "k_part_scatter":    # stable radix pass over the 8-byte symbol records
        return 2.0 * st["nsym"] * 8.0
    if kernel == "k_part_hist":
        return 1.0 * st["nsym"] * 8.0
    if kernel == "k_bucket_count":    # read every key once, write (key, count) of the solid ones; averaged over its two launches
        return (1.0 * NK * kb + st["nruns"] * (kb + 4.0)) / 2.0
    if kernel == "k_extract":         # read the bases, write one key per k-mer
        return NB + NK * kb
    if kernel == "k_chunk_ranks":     # read the 8-byte sorted record, write the 17-byte coder record; histogram row per chunk
        return st["nsym"] * (8.0 + 17.0 + 2.0)
    if kernel == "k_emit_syms":
        return st["nsym"] * 8.0 + NB
    if kernel == "k_range_encode":    # read the 17-byte record of every symbol, write the output bytes
        return st["nsym"] * 17.0 + st["dna_bytes"]
    if kernel == "k_deflate_count":   # read the quality text once (frequencies and Adler sums are small)
        return st["qual_bytes"]
    if kernel == "k_deflate_emit":    # read the quality text once, write the compressed chunk once
        return st["qual_bytes"] + st.get("qual_z_bytes", 0.0)
    if kernel == "k_deflate_gather":  # move the compressed chunks to their place
        return 2.0 * st.get("qual_z_bytes", 0.0)
    if kernel == "k_quals":           # read bases + qualities, write the smoothed text (Bloom lookups not counted)
        return 2.0 * NB + st["qual_bytes"]
    if kernel == "k_run_stats" or kernel == "k_insert_runs":
        return st["nruns"] * (8.0 + kb)
    return None


def main_b200(args):
    import torch
    import torch.distributed as dist
    import leon_b200
    from leon_b200 import synth

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

    # ---- the workload: every rank generates the same seeded read set; with N > 1 ranks the set is dealt in rounds of
    # N * S consecutive reads, S to each rank in rank order (leon_b200.shard; include/leon_b200.h leon_comm_*), so the
    # total work is fixed as N grows ("strong").
    from leon_b200 import shard
    total_reads = args.reads if args.reads else GENOME_LEN * COVERAGE // READ_LEN
    genome = synth.make_genome(GENOME_LEN, 1)
    S = 0
    if world == 1:
        pin_b = leon_b200.api.PinnedArray(total_reads * READ_LEN)
        pin_q = leon_b200.api.PinnedArray(total_reads * READ_LEN) if args.fastq else None
        bases, offs, quals = synth.sample_reads(genome, total_reads, READ_LEN, 2, out_bases=pin_b.array, with_quals=args.fastq,
                                                out_quals=pin_q.array if args.fastq else None)
    else:
        S = args.anchor_batch if args.anchor_batch else shard.choose_batch_reads(total_reads, world, RPB)
        fb, fo, fq = synth.sample_reads(genome, total_reads, READ_LEN, 2, with_quals=args.fastq)
        sb, offs, sq = shard.shard_reads(fb, fo, fq, rank, world, S)
        del fb, fq
        pin_b = leon_b200.api.PinnedArray(len(sb))
        pin_b.array[:] = sb
        bases = pin_b.array
        quals = None
        if args.fastq:
            pin_q = leon_b200.api.PinnedArray(len(sq))
            pin_q.array[:] = sq
            quals = pin_q.array
        del sb, sq
    local_reads = len(offs) - 1
    pin_o = leon_b200.api.PinnedArray((local_reads + 1) * 8, np.uint64)
    pin_o.array[:] = offs
    offs = pin_o.array

    dev_zlib = args.fastq and not args.host_zlib
    eng = leon_b200.Leon(kmer_size=KMER, reads_per_block=RPB, has_qual=args.fastq, device=local_rank,
                         anchor_batch=S if world > 1 else args.anchor_batch, qual_zlib_on_device=dev_zlib, skip_qual_text=dev_zlib)
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
    ktimes = {}
    phase = {}
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
    res = eng.fetch(copy=False)
    # the roofline describes rank 0's launches: its own share of the reads
    stats = {"nreads": local_reads, "nbases": local_reads * READ_LEN, "nkmers": local_reads * (READ_LEN - KMER + 1),
             "nruns": res["nb_distinct"], "dna_bytes": len(res["dna"]),
             "qual_bytes": float(local_reads * (READ_LEN + 1)) if args.fastq else 0.0, "qual_z_bytes": float(len(res["qual_z"]))}
    nsym = eng.debug_fetch("nsym", np.uint64)
    stats["nsym"] = float(nsym.sum())
    anchors = res["anchor_count"]
    dna_bytes = len(res["dna"])

    # ---- e2e: host buffers in, host buffers out, through the C ABI one-shot call
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
        h2d, d2h, launches = (int(x) for x in t.tolist())
    e2e_value = total_reads / (e2e_ms / 1000.0)

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
                    "share_of_step": ms_tot / dev_ms, "algorithmic_bytes_per_launch": ab}
        if dom == "k_range_encode":
            # one warp per block stream runs the serial low/range recurrence: bound by single-warp issue latency, not by
            # bytes -- the figure that describes it is cycles per symbol of the longest stream
            nblk = max(1, -(-local_reads // RPB))
            sym_per_block = stats["nsym"] / nblk
            roofline["note"] = "latency-bound serial recurrence (one warp per 50000-read block); HBM fraction is not its limit"
            roofline["ns_per_symbol"] = per_launch_ms * 1e6 / sym_per_block
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
        if ab_k and n_k:
            gbs = (ab_k / 1e9) / ((ms_tot_k / n_k) / 1e3)
            kernel_rooflines[kname] = {"ms_per_launch": ms_tot_k / n_k, "launches_per_step": n_k / args.steps,
                                       "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / peak}
    # Bloom-probe roofline of k_walk (north_star): extension steps (= contains4 windows) per second against what the
    # memory system delivers for that access pattern on a filter of this size (tools/probe_roofline.cu, measured on B200)
    probe = None
    if "k_walk" in ktimes and ktimes["k_walk"][1]:
        wms = ktimes["k_walk"][0] / ktimes["k_walk"][1]
        steps = float(stats["nbases"] - KMER * stats["nreads"])
        wps = steps / (wms / 1e3)
        probe = {"kernel": "k_walk", "windows_per_s": wps, "filter_bytes": int(len(res["bloom"])), "peak_windows_per_s": None,
                 "frac": None, "regime": None}
        try:
            pr = json.load(open(os.path.join(ROOT, "profiles", "probe_roofline.json")))["windows_per_s"]
            fb = len(res["bloom"])
            key = "l2_resident_6.9MB" if fb <= 64e6 else ("150MB" if fb <= 300e6 else ("dram_1GB" if fb <= 2e9 else "dram_4.5GB"))
            probe.update(peak_windows_per_s=pr[key], frac=wps / pr[key], regime=key, peak_source="tools/probe_roofline.cu microbenchmark")
            if wps > pr[key]:
                # the kernel issues fewer loads per window than the microbenchmark item and runs above its figure: the
                # bound that applies is the L1 request rate of the committed ncu capture of this kernel
                l1 = json.load(open(os.path.join(ROOT, "profiles", "probe_roofline.json")))["k_walk_ncu"]["l1tex_throughput_pct"] / 100.0
                probe.update(peak_windows_per_s=wps / l1, frac=l1, microbenchmark_windows_per_s=pr[key],
                             peak_source="ncu l1tex__throughput of k_walk (profiles/r1i_ncu_full.csv); above the microbenchmark figure")
        except Exception:
            pass
    if args.phases and rank == 0:
        print("phases (ms/step):", {k: round(v / args.steps, 3) for k, v in phase.items()}, file=sys.stderr)
        print("kernels (ms/step, launches/step):",
              {k: (round(v[0] / args.steps, 3), v[1] // args.steps) for k, v in sorted(ktimes.items(), key=lambda kv: -kv[1][0])},
              file=sys.stderr)
        print("anchors %d, dna bytes %d, symbols %d, solid %d, cutoff %d" % (
            anchors, dna_bytes, int(stats["nsym"]), res["nb_solids_cutoff"], res["cutoff"]), file=sys.stderr)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_block()
    eng.close()
    if rank != 0:
        return 0
    line = {
        "metric": METRIC, "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic",
        "config": {"workload": (WORKLOAD if not args.fastq else "DEBUG (configs[2] shape, FASTQ lossy)") if not args.reads else "DEBUG SIZE (invalid as a bench line): %d reads" % total_reads,
                   "reads": total_reads, "read_len": READ_LEN, "kmer_size": KMER, "reads_per_block": RPB,
                   "l2": "inputs larger than L2 (460 MB of bases, 2.6 GB of keys per step); no flush needed",
                   "parallelism": "1 process per GPU" if world == 1 else
                   "%d ranks, reads dealt in rounds of %d x %d; NCCL: k-mer all-to-all by hash owner, Bloom all-gather + OR, "
                   "anchor candidates all-gathered per batch" % (world, world, S)},
        "input_gbs": value * READ_LEN / 1e9,
        "e2e": {"value": e2e_value, "unit": "reads/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms, "input_gbs": e2e_value * READ_LEN / 1e9},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "kernel_rooflines": kernel_rooflines,
        "probe_roofline": probe,
        "cpu_baseline": cpu,
        "phases_ms_per_step": {k: v / args.steps for k, v in phase.items()},
        "kernels_ms_per_step": {k: v[0] / args.steps for k, v in ktimes.items()},
    }
    print(json.dumps(line))
    return 0


if __name__ == "__main__":
    a = parse_args()
    sys.exit(main_reference(a) if a.impl == "reference" else main_b200(a))
