#!/usr/bin/env bash
mkdir -p gpurun_out
for conn in 8 32; do
  CUDA_DEVICE_MAX_CONNECTIONS=$conn timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/r2m_conn$conn.json 2>/dev/null
  python - $conn <<'PY'
import json,sys
b=json.load(open('gpurun_out/r2m_conn%s.json'%sys.argv[1])); p=b['phases_ms_per_step']; l=b['large']
print("connections %-3s c2 %.1f ms (anchors %.1f tail %.1f) e2e %.1f | c1 %.1f | large %.0f ms %s | rc %.1f ms/launch" % (sys.argv[1], b['ms_per_step'], p['anchor_ms'], p['walk_ms'], b['e2e']['ms_per_step'], b['configs1']['ms_per_step'], l['ms_per_step'], l['phases_s_rank0'], b['roofline']['launch_ms']))
PY
done
