#!/usr/bin/env bash
# round 2 evidence pack: independent probe roofline, launch list + ncu --set full of the bench command (configs[2]),
# ncu --set full of k_walk / k_quals at configs[3] scale (150 MB filter, DRAM regime)
set -x
mkdir -p gpurun_out
./tools/bin/probe_roofline > gpurun_out/r2_probe_roofline.txt 2>&1
tail -3 gpurun_out/r2_probe_roofline.txt
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extra --no-large"
$CMD > gpurun_out/r2h_plain.json 2> gpurun_out/r2h_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2h_launches.csv $CMD > gpurun_out/r2h_ncu_launches.log 2>&1
echo "launch list rc $?"
ncu --set full --clock-control none --import-source on -k 'regex:k_walk|k_range_encode|k_msd_scatter|k_msd_hist|k_bucket_count|k_quals|k_extract|k_chunk_ranks|k_anchor_lookup|k_deflate_emit' \
    -s 60 -c 24 -f -o gpurun_out/r2h_full $CMD > gpurun_out/r2h_ncu_full.log 2>&1
echo "full rc $?"
LCMD="python bench.py --only-large --steps 1 --warmup 1"
$LCMD > gpurun_out/r2h_large_plain.json 2> gpurun_out/r2h_large_plain.err &&
ncu --set full --clock-control none --import-source on -k 'regex:k_walk|k_quals|k_anchor_lookup|k_candidates' -s 40 -c 8 -f -o gpurun_out/r2h_full_c3 $LCMD > gpurun_out/r2h_ncu_full_c3.log 2>&1
echo "full c3 rc $?"
ls -la gpurun_out | grep r2h
