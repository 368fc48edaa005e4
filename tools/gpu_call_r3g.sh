#!/usr/bin/env bash
# multiplicative hash of the anchor dictionary: full GPU suite, bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r3g_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r3g_pytest_gpu.log
B="python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline --no-extra --no-large"
timeout 600 $B > gpurun_out/r3g_bench.json 2> gpurun_out/r3g_bench_phases.txt; echo "bench rc $?"
python - <<'PY'
import json
b = [json.loads(l) for l in open('gpurun_out/r3g_bench.json') if l.startswith('{')][-1]
print(round(b['ms_per_step'], 2), 'e2e', round(b['e2e']['ms_per_step'], 2), b['digest_ok'], {x: round(v, 2) for x, v in b['phases_ms_per_step'].items() if x in ('count_ms', 'sort_ms', 'anchor_ms', 'walk_ms', 'total_ms')})
PY
