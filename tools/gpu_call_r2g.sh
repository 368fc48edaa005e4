#!/usr/bin/env bash
# pipelining experiments on configs[1] / configs[2]: deferred (default) vs pipelined, with and without a shared-memory claim
# that keeps k_walk at 3 / 2 CTAs per SM
mkdir -p gpurun_out
run() { # tag, env...
  tag=$1; shift
  env "$@" timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-large > gpurun_out/r2g_$tag.json 2>/dev/null
  python - "$tag" <<'PY'
import json,sys
b=json.load(open('gpurun_out/r2g_%s.json'%sys.argv[1]))
p=b['phases_ms_per_step']; q=b['configs1']['phases_ms_per_step']
print("%-22s c2 %.1f ms (anchors %.1f tail %.1f) e2e %.1f | c1 %.1f ms (anchors %.1f tail %.1f) digest %s %s" % (sys.argv[1], b['ms_per_step'], p['anchor_ms'], p['walk_ms'], b['e2e']['ms_per_step'], b['configs1']['ms_per_step'], q['anchor_ms'], q['walk_ms'], b['digest_ok'], b['configs1']['digest_ok']))
PY
}
run deferred X=1
run pipelined LEON_B200_SB_PIPELINE_MIN=0
run pipelined_smem76k LEON_B200_SB_PIPELINE_MIN=0 LEON_B200_WALK_SMEM=76000
run pipelined_smem114k LEON_B200_SB_PIPELINE_MIN=0 LEON_B200_WALK_SMEM=114000
run deferred_smem76k LEON_B200_WALK_SMEM=76000
