#!/usr/bin/env bash
# 4 tag probes in flight in k_anchor_lookup; two walk lanes (LEON_B200_WALK_LANES=1|2): suite, bench A/B, streamed configs[3]
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r3b_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r3b_pytest_gpu.log
B="python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline --no-extra --no-large"
for L in 1 2; do
  LEON_B200_WALK_LANES=$L timeout 600 $B > gpurun_out/r3b_bench_lanes$L.json 2> gpurun_out/r3b_bench_lanes${L}_phases.txt; echo "bench lanes $L rc $?"
done
python - <<'PY'
import json
for n in (1, 2):
    b = [json.loads(l) for l in open('gpurun_out/r3b_bench_lanes%d.json' % n) if l.startswith('{')][-1]
    print('lanes', n, round(b['ms_per_step'], 2), 'e2e', round(b['e2e']['ms_per_step'], 2), b['digest_ok'], {k: round(v, 2) for k, v in b['phases_ms_per_step'].items() if k in ('count_ms', 'anchor_ms', 'walk_ms', 'total_ms')}, 'k_walk', round(b['kernels_ms_per_step']['k_walk'], 2), 'rc', round(b['roofline']['cycles_per_symbol'], 1))
PY
for L in 1 2; do
LEON_B200_WALK_LANES=$L timeout 900 python bench.py --steps 2 --warmup 2 --only-large > gpurun_out/r3b_large_lanes$L.json 2> gpurun_out/r3b_large_lanes$L.err; echo "large lanes $L rc $?"
python - <<PY
import json
b = [json.loads(l) for l in open('gpurun_out/r3b_large_lanes$L.json') if l.startswith('{')][-1]
l = b.get('large', b)
print($L, {k: l.get(k) for k in ('value', 'ms_per_step', 'digest_ok')}, {k: round(l['phases_s_rank0'][k], 3) for k in ('count_s', 'encode_s')}, [x['call_ms'] for x in l['phases_s_rank0']['batches_ms']])
PY
done
