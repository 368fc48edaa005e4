#!/usr/bin/env bash
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench_phases.txt; echo "bench rc $?"
tail -4 gpurun_out/r2c_bench_phases.txt
python - <<'PY'
import json
b=json.load(open('gpurun_out/r2c_bench.json'))
print({k:b[k] for k in ('value','ms_per_step','digest_ok')}, b['e2e']['ms_per_step'], b['configs1']['ms_per_step'], b['configs1']['phases_ms_per_step'])
PY
