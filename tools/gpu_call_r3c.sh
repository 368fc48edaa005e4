#!/usr/bin/env bash
# k_msd_scatter with 512 threads (8 keys per thread, 48 warps per SM) against 256 (16 keys, 32 warps)
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline --no-extra --no-large"
for T in 256 512; do
  LEON_B200_MSD_THREADS=$T timeout 600 $B > gpurun_out/r3c_bench_msd$T.json 2> gpurun_out/r3c_bench_msd${T}_phases.txt; echo "bench msd $T rc $?"
done
python - <<'PY'
import json
for n in (256, 512):
    b = [json.loads(l) for l in open('gpurun_out/r3c_bench_msd%d.json' % n) if l.startswith('{')][-1]
    print('msd threads', n, round(b['ms_per_step'], 2), 'e2e', round(b['e2e']['ms_per_step'], 2), b['digest_ok'], {k: round(v, 2) for k, v in b['phases_ms_per_step'].items() if k in ('count_ms', 'sort_ms', 'anchor_ms', 'walk_ms', 'total_ms')}, 'k_msd_scatter', round(b['kernels_ms_per_step']['k_msd_scatter'], 3))
PY
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or seeded or k63 or stream" > gpurun_out/r3c_pytest.log 2>&1; tail -2 gpurun_out/r3c_pytest.log
LEON_B200_MSD_THREADS=512 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or seeded or stream" > gpurun_out/r3c_pytest512.log 2>&1; tail -2 gpurun_out/r3c_pytest512.log
