#!/usr/bin/env bash
# dictionary tags, offsets-first uploads, batch 0 announced behind the counting pass: full GPU suite, bench, streamed configs[3]
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2y_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r2y_pytest_gpu.log
B="python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline --no-extra --no-large"
timeout 600 $B > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench_phases.txt; echo "bench rc $?"
timeout 900 python bench.py --steps 2 --warmup 3 --only-large > gpurun_out/r2y_large.json 2> gpurun_out/r2y_large.err; echo "large rc $?"
LEON_B200_TRACE=1 timeout 900 python bench.py --steps 1 --warmup 3 --only-large > gpurun_out/r2y_large_trace.json 2> gpurun_out/r2y_large_trace.err; echo "trace rc $?"
grep "trace" gpurun_out/r2y_large_trace.err | tail -24
python - <<'PY'
import json
b = [json.loads(l) for l in open('gpurun_out/r2y_bench.json') if l.startswith('{')][-1]
print(round(b['ms_per_step'], 2), 'e2e', round(b['e2e']['ms_per_step'], 2), b['digest_ok'], {k: round(v, 2) for k, v in b['phases_ms_per_step'].items() if k in ('count_ms', 'anchor_ms', 'walk_ms', 'total_ms')})
for n in ('large', 'large_trace'):
    try:
        b = [json.loads(l) for l in open('gpurun_out/r2y_%s.json' % n) if l.startswith('{')][-1]
        l = b.get('large', b)
        print(n, {k: l.get(k) for k in ('value', 'ms_per_step', 'count_passes', 'digest_ok', 'digest_mismatch')})
        p = l.get('phases_s_rank0'); print({k: p[k] for k in ('count_s', 'bloom_s', 'encode_s', 'batches_ms')})
    except Exception as e:
        print(n, 'failed', e)
PY
