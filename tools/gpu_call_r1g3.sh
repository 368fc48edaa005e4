#!/usr/bin/env bash
# Third GPU-box visit of r1g: Bloom probe loads with an L2 cache hint (now the default).  Microbenchmark of the load
# modes, the whole GPU suite with the hinted loads, bench lines with the hint on / off (FASTA and FASTQ), the
# configs[3]-scale run with an evict_normal policy (is it the policy or the load path?), and k_walk's memory metrics.
set -u
mkdir -p gpurun_out
tools/bin/probe_roofline > gpurun_out/probe_roofline_modes.txt 2>&1; echo "probe rc=$?"; cat gpurun_out/probe_roofline_modes.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 5 --warmup 3 --phases > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
LEON_B200_L2_KEEP=0 python bench.py --steps 5 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err; echo "bench plain rc=$?"
python bench.py --fastq --steps 5 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_fastq.json 2> gpurun_out/bench_fastq.err; echo "bench fastq rc=$?"
LEON_B200_L2_KEEP=0 python bench.py --fastq --steps 5 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_fastq_plain.json 2> gpurun_out/bench_fastq_plain.err; echo "bench fastq plain rc=$?"
grep -h -o "ms_per_step\": [0-9.]*" gpurun_out/bench.json gpurun_out/bench_plain.json gpurun_out/bench_fastq.json gpurun_out/bench_fastq_plain.json
grep -h -o "'k_walk': ([0-9.]*\|'k_quals': ([0-9.]*" gpurun_out/bench.err gpurun_out/bench_plain.err gpurun_out/bench_fastq.err gpurun_out/bench_fastq_plain.err
LEON_B200_L2_KEEP=2 timeout 600 python tools/run_scale.py --genome 100000000 --reads 33333333 --len 150 > gpurun_out/c4_hint_normal.log 2>&1; echo "c4 evict_normal rc=$?"
grep -h -o '"k_walk": [0-9.]*\|digest.*\|"wall_s": [0-9.]*' gpurun_out/c4_hint_normal.log
M="lts__t_sector_hit_rate.pct,lts__t_sectors_op_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active"
timeout 500 ncu --metrics $M --clock-control none -k regex:k_walk -c 1 --csv --log-file gpurun_out/c4_walk_default.csv \
    python tools/run_scale.py --genome 100000000 --reads 13333333 --len 150 --runs 1 > gpurun_out/c4_ncu_default.log 2>&1; echo "ncu walk default rc=$?"
LEON_B200_L2_KEEP=0 timeout 500 ncu --metrics $M --clock-control none -k regex:k_walk -c 1 --csv --log-file gpurun_out/c4_walk_plain.csv \
    python tools/run_scale.py --genome 100000000 --reads 13333333 --len 150 --runs 1 > gpurun_out/c4_ncu_plain.log 2>&1; echo "ncu walk plain rc=$?"
