#!/usr/bin/env bash
# host-clock trace of the streamed configs[3] run; pipelined stage for 16 super-blocks (LEON_B200_SB_PIPELINE_MIN=8) A/B
mkdir -p gpurun_out
LEON_B200_TRACE=1 timeout 900 python bench.py --steps 1 --warmup 3 --only-large > gpurun_out/r2w_large_trace.json 2> gpurun_out/r2w_large_trace.err; echo "trace rc $?"
grep "trace" gpurun_out/r2w_large_trace.err | tail -24
LEON_B200_SB_PIPELINE_MIN=8 timeout 900 python bench.py --steps 2 --warmup 3 --only-large > gpurun_out/r2w_large_pipe.json 2> gpurun_out/r2w_large_pipe.err; echo "pipe rc $?"
python - <<'PY'
import json
for n in ('large_trace', 'large_pipe'):
    try:
        b = [json.loads(l) for l in open('gpurun_out/r2w_%s.json' % n) if l.startswith('{')][-1]
        l = b.get('large', b)
        print(n, {k: l.get(k) for k in ('value', 'ms_per_step', 'count_passes', 'digest_ok', 'digest_mismatch')})
        p = l.get('phases_s_rank0'); print({k: p[k] for k in ('count_s', 'bloom_s', 'encode_s', 'batches_ms')})
    except Exception as e:
        print(n, 'failed', e)
PY
