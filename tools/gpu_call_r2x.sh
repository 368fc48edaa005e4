#!/usr/bin/env bash
# queued prefetch: streaming tests, then the streamed configs[3] run with the host-clock trace and without
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_file.py -m gpu -x -q -k "stream or Stream" > gpurun_out/r2x_pytest_stream.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r2x_pytest_stream.log
timeout 900 python bench.py --steps 2 --warmup 3 --only-large > gpurun_out/r2x_large.json 2> gpurun_out/r2x_large.err; echo "large rc $?"
LEON_B200_TRACE=1 timeout 900 python bench.py --steps 1 --warmup 3 --only-large > gpurun_out/r2x_large_trace.json 2> gpurun_out/r2x_large_trace.err; echo "trace rc $?"
grep "trace" gpurun_out/r2x_large_trace.err | grep -v "anchors:" | tail -22
python - <<'PY'
import json
for n in ('large', 'large_trace'):
    try:
        b = [json.loads(l) for l in open('gpurun_out/r2x_%s.json' % n) if l.startswith('{')][-1]
        l = b.get('large', b)
        print(n, {k: l.get(k) for k in ('value', 'ms_per_step', 'count_passes', 'digest_ok', 'digest_mismatch')})
        p = l.get('phases_s_rank0'); print({k: p[k] for k in ('count_s', 'bloom_s', 'encode_s', 'batches_ms')})
    except Exception as e:
        print(n, 'failed', e)
PY
