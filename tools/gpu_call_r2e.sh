#!/usr/bin/env bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2e_pytest_gpu.log
tail -5 gpurun_out/r2e_pytest_gpu.log
timeout 1200 python bench.py --steps 10 --warmup 3 --phases > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench_phases.txt; echo "bench rc $?"
tail -5 gpurun_out/r2e_bench_phases.txt
python - <<'PY'
import json
b=json.load(open('gpurun_out/r2e_bench.json'))
print({k:b[k] for k in ('value','ms_per_step','digest_ok')}, b['e2e']['ms_per_step'], b['configs1']['ms_per_step'])
print(json.dumps(b.get('large'))[:1800])
PY
