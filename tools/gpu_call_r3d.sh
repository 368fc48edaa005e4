#!/usr/bin/env bash
# k_msd_scatter 512 against 1024 threads; k_bucket_count with a plain load before the compare-and-swap (both runs)
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline --no-extra --no-large"
for T in 512 1024; do
  LEON_B200_MSD_THREADS=$T timeout 600 $B > gpurun_out/r3d_bench_msd$T.json 2> gpurun_out/r3d_bench_msd${T}_phases.txt; echo "bench msd $T rc $?"
done
python - <<'PY'
import json
for n in (512, 1024):
    b = [json.loads(l) for l in open('gpurun_out/r3d_bench_msd%d.json' % n) if l.startswith('{')][-1]
    k = b['kernels_ms_per_step']
    print('msd threads', n, round(b['ms_per_step'], 2), 'e2e', round(b['e2e']['ms_per_step'], 2), b['digest_ok'], {x: round(v, 2) for x, v in b['phases_ms_per_step'].items() if x in ('count_ms', 'sort_ms', 'anchor_ms', 'walk_ms', 'total_ms')}, 'k_msd_scatter', round(k['k_msd_scatter'], 3), 'k_bucket_count', round(k['k_bucket_count'], 3), 'k_msd_hist', round(k['k_msd_hist'], 3))
PY
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or seeded or stream or edge" > gpurun_out/r3d_pytest.log 2>&1; tail -2 gpurun_out/r3d_pytest.log
LEON_B200_MSD_THREADS=1024 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or seeded or stream" > gpurun_out/r3d_pytest1024.log 2>&1; tail -2 gpurun_out/r3d_pytest1024.log
