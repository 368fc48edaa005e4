#!/usr/bin/env bash
# quality kernels in the idle tail of the range coder chains (LEON_B200_QUALS_LATE A/B), count budget 50 % on configs[3]
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --phases --no-cpu-baseline --no-extra --no-large"
LEON_B200_QUALS_LATE=0 timeout 600 $B > gpurun_out/r2t_early.json 2> gpurun_out/r2t_early_phases.txt; echo "early rc $?"
LEON_B200_QUALS_LATE=1 timeout 600 $B > gpurun_out/r2t_late.json 2> gpurun_out/r2t_late_phases.txt; echo "late rc $?"
python - <<'PY'
import json
for n in ('early', 'late'):
    try:
        b = [json.loads(l) for l in open('gpurun_out/r2t_%s.json' % n) if l.startswith('{')][-1]
        print(n, round(b['ms_per_step'], 2), 'e2e', round(b['e2e']['ms_per_step'], 2), b['digest_ok'],
              {k: round(v, 2) for k, v in b['phases_ms_per_step'].items() if k in ('count_ms', 'anchor_ms', 'walk_ms', 'total_ms')},
              'k_quals', b['kernels_ms_per_step'].get('k_quals'), 'rc', b['roofline'].get('cycles_per_symbol'))
    except Exception as e:
        print(n, 'failed', e)
PY
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2t_pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r2t_pytest_gpu.log
timeout 900 python bench.py --steps 2 --warmup 3 --only-large > gpurun_out/r2t_large.json 2> gpurun_out/r2t_large.err; echo "large rc $?"
python - <<'PY'
import json
try:
    b = [json.loads(l) for l in open('gpurun_out/r2t_large.json') if l.startswith('{')][-1]
    l = b.get('large', b)
    print({k: l.get(k) for k in ('value', 'ms_per_step', 'count_passes', 'peak_device_bytes_per_rank', 'digest_ok', 'digest_mismatch')})
    print(l.get('phases_s_rank0'))
except Exception as e:
    print('large failed', e)
PY
