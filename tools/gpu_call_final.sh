#!/usr/bin/env bash
# One GPU-box visit for the state of the tree at round end: parity tests, smoke, both bench arms, FASTQ bench line,
# launch lists and ncu --set full captures of the main kernels (FASTA run) and of the quality kernels (FASTQ run),
# file-level benchmark.  Every ncu run comes after the same command has exited 0 without ncu.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 5 --warmup 3 --phases > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python bench.py --fastq --steps 5 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_fastq.json 2> gpurun_out/bench_fastq.err; echo "bench fastq rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_fastq.csv \
    python bench.py --fastq --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches_fastq.log 2>&1; echo "ncu launches fastq rc=$?"
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_range_encode|k_walk|k_msd_scatter|k_msd_hist|k_bucket_count|k_chunk_ranks|k_extract|k_anchor_lookup|k_part_scatter' -c 12 -f -o gpurun_out/full \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_quals|k_deflate_count|k_deflate_codes|k_deflate_emit|k_deflate_gather' -c 6 -f -o gpurun_out/full_fastq \
    python bench.py --fastq --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_full_fastq.log 2>&1; echo "ncu full fastq rc=$?"
python tools/file_bench.py --inproc 3 --out gpurun_out/file_bench_fasta.json > gpurun_out/file_bench_fasta.log 2>&1; echo "file bench fasta rc=$?"
python tools/file_bench.py --fastq --inproc 3 --out gpurun_out/file_bench_fastq.json > gpurun_out/file_bench_fastq.log 2>&1; echo "file bench fastq rc=$?"
grep -h "in-process\|^ours run\|^reference\|reference decode" gpurun_out/file_bench_fasta.log gpurun_out/file_bench_fastq.log
cut -c1-300 gpurun_out/bench.json; cut -c1-300 gpurun_out/bench_fastq.json
