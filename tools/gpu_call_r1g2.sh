#!/usr/bin/env bash
# Second GPU-box visit of r1g: (1) configs[3]-scale run (150 MB filter) with the default k_walk and with the L2
# evict_last variant, plus the L2 hit rate / DRAM bytes of k_walk from ncu at 20x coverage of the same genome;
# (2) GPU decoder: many-block case (920 blocks) and an ncu --set full capture of the kept kernel; (3) ncu --set full
# of the quality kernels of the FASTQ run.  Every ncu run comes after the same command has exited 0 without ncu.
set -u
mkdir -p gpurun_out
M="lts__t_sector_hit_rate.pct,lts__t_sector_op_read_hit_rate.pct,lts__t_sectors_op_read.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed"
timeout 600 python tools/run_scale.py --genome 100000000 --reads 33333333 --len 150 > gpurun_out/c4_base.log 2>&1; echo "c4 base rc=$?"
LEON_B200_WALK_KEEP=1 timeout 600 python tools/run_scale.py --genome 100000000 --reads 33333333 --len 150 > gpurun_out/c4_keep.log 2>&1; echo "c4 keep rc=$?"
grep -h -o '"k_walk": [0-9.]*\|digest.*' gpurun_out/c4_base.log gpurun_out/c4_keep.log
timeout 500 ncu --metrics $M --clock-control none -k regex:k_walk -c 1 --csv --log-file gpurun_out/c4_walk_base.csv \
    python tools/run_scale.py --genome 100000000 --reads 13333333 --len 150 --runs 1 > gpurun_out/c4_ncu_base.log 2>&1; echo "ncu walk base rc=$?"
LEON_B200_WALK_KEEP=1 timeout 500 ncu --metrics $M --clock-control none -k regex:k_walk -c 1 --csv --log-file gpurun_out/c4_walk_keep.csv \
    python tools/run_scale.py --genome 100000000 --reads 13333333 --len 150 --runs 1 > gpurun_out/c4_ncu_keep.log 2>&1; echo "ncu walk keep rc=$?"
tail -n 8 gpurun_out/c4_walk_base.csv | cut -c1-300; tail -n 8 gpurun_out/c4_walk_keep.csv | cut -c1-300
timeout 600 python tools/decode_bench.py --rpb 5000 --out gpurun_out/decode_bench_920.json > gpurun_out/decode_bench_920.log 2>&1; echo "decode 920 rc=$?"; cat gpurun_out/decode_bench_920.json
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:k_decode_blocks' -c 1 -f -o gpurun_out/full_decode_v1 \
    python tools/decode_bench.py --reads 460000 --rpb 5000 > gpurun_out/ncu_full_decode_v1.log 2>&1; echo "ncu full decode rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on \
    -k 'regex:k_quals|k_deflate_count|k_deflate_codes|k_deflate_emit|k_deflate_gather' -c 6 -f -o gpurun_out/full_fastq \
    python bench.py --fastq --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_full_fastq.log 2>&1; echo "ncu full fastq rc=$?"
