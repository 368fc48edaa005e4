#!/usr/bin/env bash
# exclusive SMs for the late chains (LEON_B200_RC_EXCLUSIVE = number of last super-blocks), per-batch phases of configs[3]
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline --no-extra --no-large"
for P in 0 3 5 10; do
  LEON_B200_RC_EXCLUSIVE=$P timeout 600 $B > gpurun_out/r2u_excl$P.json 2> gpurun_out/r2u_excl${P}_phases.txt; echo "excl $P rc $?"
done
python - <<'PY'
import json
for n in (0, 3, 5, 10):
    try:
        b = [json.loads(l) for l in open('gpurun_out/r2u_excl%d.json' % n) if l.startswith('{')][-1]
        print(n, round(b['ms_per_step'], 2), 'e2e', round(b['e2e']['ms_per_step'], 2), b['digest_ok'],
              {k: round(v, 2) for k, v in b['phases_ms_per_step'].items() if k in ('count_ms', 'anchor_ms', 'walk_ms', 'total_ms')},
              'k_quals', round(b['kernels_ms_per_step'].get('k_quals', 0), 2), 'k_walk', round(b['kernels_ms_per_step'].get('k_walk', 0), 2), 'rc', round(b['roofline'].get('cycles_per_symbol'), 1))
    except Exception as e:
        print(n, 'failed', e)
PY
timeout 900 python bench.py --steps 2 --warmup 3 --only-large > gpurun_out/r2u_large.json 2> gpurun_out/r2u_large.err; echo "large rc $?"
python - <<'PY'
import json
try:
    b = [json.loads(l) for l in open('gpurun_out/r2u_large.json') if l.startswith('{')][-1]
    l = b.get('large', b)
    print({k: l.get(k) for k in ('value', 'ms_per_step', 'count_passes', 'peak_device_bytes_per_rank', 'digest_ok', 'digest_mismatch')})
    print(l.get('phases_s_rank0'))
except Exception as e:
    print('large failed', e)
PY
