#!/usr/bin/env bash
# final evidence of this tree: GPU suite, the driver's two bench commands, launch list and ncu --set full of the bench command
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r3e_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/r3e_pytest_gpu.log
tail -4 gpurun_out/r3e_pytest_gpu.log
python __graft_entry__.py smoke 2>&1 | tail -1
timeout 1500 python bench.py --steps 10 --warmup 3 --phases > gpurun_out/r3e_bench.json 2> gpurun_out/r3e_bench_phases.txt; echo "bench rc $?"
python - <<'PY'
import json
b=json.load(open('gpurun_out/r3e_bench.json'))
print({k:b[k] for k in ('value','ms_per_step','digest_ok')}, 'e2e', b['e2e']['ms_per_step'], 'c1', b['configs1']['ms_per_step'], b['configs1']['digest_ok'])
print(b['cpu_baseline']['value'], {k:(round(v,3) if isinstance(v,float) else v) for k,v in b['file_e2e'].items() if k!='what'})
l=b['large']; print({k:l[k] for k in ('value','ms_per_step','count_passes','batches_per_rank','phases_s_rank0','peak_device_bytes_per_rank','digest_ok')})
print(b['roofline']); print(b['probe_roofline'])
PY
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extra --no-large"
$CMD > gpurun_out/r3e_plain.json 2> gpurun_out/r3e_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r3e_launches.csv $CMD > gpurun_out/r3e_ncu_launches.log 2>&1
echo "launch list rc $?"
ncu --set full --clock-control none --import-source on -k 'regex:k_walk|k_range_encode|k_extract|k_msd_scatter|k_msd_hist|k_bucket_count|k_quals|k_anchor_lookup' -c 16 -f -o gpurun_out/r3e_full $CMD > gpurun_out/r3e_ncu_full.log 2>&1
echo "full rc $?"
