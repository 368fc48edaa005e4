#!/usr/bin/env bash
# Fourth GPU-box visit of r1g: k_walk with all 14 probe loads of a step in flight (the fix the load-mode experiment led
# to).  GPU suite, smoke, both bench arms, FASTQ line, configs[3]-scale run, launch list, ncu --set full of the main
# kernels of the bench command and k_walk's memory metrics at configs[3] scale (20x coverage of the same genome).
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 5 --warmup 3 --phases > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python bench.py --fastq --steps 5 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_fastq.json 2> gpurun_out/bench_fastq.err; echo "bench fastq rc=$?"
grep -h -o "ms_per_step\": [0-9.]*" gpurun_out/bench.json gpurun_out/bench_fastq.json
grep -h -o "'k_walk': ([0-9.]*\|'k_quals': ([0-9.]*" gpurun_out/bench.err gpurun_out/bench_fastq.err
timeout 600 python tools/run_scale.py --genome 100000000 --reads 33333333 --len 150 > gpurun_out/c4_run.log 2>&1; echo "c4 rc=$?"
grep -h -o '"k_walk": [0-9.]*\|digest.*\|"wall_s": [0-9.]*' gpurun_out/c4_run.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_range_encode|k_walk|k_msd_scatter|k_msd_hist|k_bucket_count|k_chunk_ranks|k_extract|k_anchor_lookup|k_part_scatter' -c 12 -f -o gpurun_out/full \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
M="lts__t_sector_hit_rate.pct,lts__t_sectors_op_read.sum,l1tex__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active"
timeout 500 ncu --metrics $M --clock-control none -k regex:k_walk -c 1 --csv --log-file gpurun_out/c4_walk_final.csv \
    python tools/run_scale.py --genome 100000000 --reads 13333333 --len 150 --runs 1 > gpurun_out/c4_ncu_final.log 2>&1; echo "ncu walk c4 rc=$?"
tail -n 12 gpurun_out/c4_walk_final.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}'
