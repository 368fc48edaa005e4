#!/usr/bin/env bash
# pipelined stage: quality kernels started when 75 % / 60 % / 90 % of the super-blocks are walked, against at once (0)
mkdir -p gpurun_out
for P in 0 75 60 90; do
  LEON_B200_QUALS_LATE_PIPE=$P timeout 600 python bench.py --steps 2 --warmup 2 --only-large > gpurun_out/r3a_large_pipe$P.json 2> gpurun_out/r3a_large_pipe$P.err; echo "pipe $P rc $?"
  python - <<PY
import json
b = [json.loads(l) for l in open('gpurun_out/r3a_large_pipe$P.json') if l.startswith('{')][-1]
l = b.get('large', b)
print($P, {k: l.get(k) for k in ('value', 'ms_per_step', 'digest_ok')}, {k: round(l['phases_s_rank0'][k], 3) for k in ('count_s', 'encode_s')}, [x['call_ms'] for x in l['phases_s_rank0']['batches_ms']])
PY
done
