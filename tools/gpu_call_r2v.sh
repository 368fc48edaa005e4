#!/usr/bin/env bash
# ragged k_extract (per-warp transposition), exclusive SMs default 5, anchor stage trace at configs[3] scale
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2v_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r2v_pytest_gpu.log
B="python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline --no-extra --no-large"
timeout 600 $B > gpurun_out/r2v_bench.json 2> gpurun_out/r2v_bench_phases.txt; echo "bench rc $?"
timeout 900 python bench.py --steps 2 --warmup 3 --only-large > gpurun_out/r2v_large.json 2> gpurun_out/r2v_large.err; echo "large rc $?"
LEON_B200_TRACE=1 timeout 900 python bench.py --steps 1 --warmup 3 --only-large > gpurun_out/r2v_large_trace.json 2> gpurun_out/r2v_large_trace.err; echo "trace rc $?"
python - <<'PY'
import json
b = [json.loads(l) for l in open('gpurun_out/r2v_bench.json') if l.startswith('{')][-1]
print(round(b['ms_per_step'], 2), 'e2e', round(b['e2e']['ms_per_step'], 2), b['digest_ok'], b['phases_ms_per_step'])
for n in ('large', 'large_trace'):
    try:
        b = [json.loads(l) for l in open('gpurun_out/r2v_%s.json' % n) if l.startswith('{')][-1]
        l = b.get('large', b)
        print(n, {k: l.get(k) for k in ('value', 'ms_per_step', 'count_passes', 'digest_ok', 'digest_mismatch')})
        print(l.get('phases_s_rank0'))
    except Exception as e:
        print(n, 'failed', e)
PY
grep "trace" gpurun_out/r2v_large_trace.err | tail -12
