#!/usr/bin/env bash
# 4 GPUs: parity of the multi-GPU path against the oracle, then the anchor batch size (reads per rank and round) sweep
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29519"
timeout 900 $TR tools/multigpu_check.py > gpurun_out/r2k_multigpu_check_4gpu.log 2>&1; echo "check rc $?"
grep "^case\|ALL OK\|FAILED" gpurun_out/r2k_multigpu_check_4gpu.log
for S in 0 50000 100000 200000; do
  timeout 600 $TR bench.py --gpus 4 --steps 5 --warmup 3 --no-large --no-extra --anchor-batch $S > gpurun_out/r2k_bench_4gpu_S$S.json 2> gpurun_out/r2k_err_S$S.txt
  python - $S <<'PY'
import json,sys
txt=[l for l in open('gpurun_out/r2k_bench_4gpu_S%s.json'%sys.argv[1]) if l.startswith('{')]
b=json.loads(txt[-1]); p=b['phases_ms_per_step']
print("S=%-7s  %.1f ms (count %.1f anchors %.1f [batches %.0f rounds %.0f] tail %.1f dict_host %.1f) e2e %.1f digest %s" % (sys.argv[1], b['ms_per_step'], p['count_ms'], p['anchor_ms'], p['anchor_batches'], p['anchor_rounds'], p['walk_ms'], p['dict_host_ms'], b['e2e']['ms_per_step'], b['digest_ok']))
PY
done
