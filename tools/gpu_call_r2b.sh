#!/usr/bin/env bash
# round 2, GPU visit b: pipelined walk / entropy stage -- parity suite + bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2b_pytest_gpu.log
tail -15 gpurun_out/r2b_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench_phases.txt; echo "bench rc $?"
cat gpurun_out/r2b_bench_phases.txt | tail -5
python - <<'PY'
import json
b=json.load(open('gpurun_out/r2b_bench.json'))
print({k:b[k] for k in ('value','ms_per_step','digest_ok','digest_mismatch')}, b['e2e'], b.get('configs1'))
PY
