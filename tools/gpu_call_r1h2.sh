#!/usr/bin/env bash
# configs[3]-scale FASTQ run on one GPU (lossy qualities: k_quals in the DRAM regime), then ncu --set full of k_walk and
# k_range_encode of the bench command (the r1h capture stopped before them).
set -u
mkdir -p gpurun_out
timeout 1200 python tools/run_scale.py --genome 100000000 --reads 33333333 --len 150 --fastq > gpurun_out/c4_fastq_run.log 2>&1; echo "c4 fastq rc=$?"
cut -c1-700 gpurun_out/c4_fastq_run.log | tail -5
ncu --set full --clock-control none --import-source on -k 'regex:k_walk|k_range_encode$|k_chunk_ranks|k_emit_syms' -c 4 -f -o gpurun_out/full_walk \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_full_walk.log 2>&1; echo "ncu full walk rc=$?"
