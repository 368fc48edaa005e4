#!/usr/bin/env bash
# 2 GPUs: byte parity of the multi-GPU path (one-shot and streaming) against the oracle, bench lines at N=2
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $TR tools/multigpu_check.py > gpurun_out/r2f_multigpu_check_2gpu.log 2>&1; echo "check rc $?"
tail -12 gpurun_out/r2f_multigpu_check_2gpu.log
timeout 900 $TR bench.py --gpus 2 --steps 5 --warmup 3 --no-large --phases > gpurun_out/r2f_bench_2gpu.json 2> gpurun_out/r2f_bench_2gpu_phases.txt; echo "bench rc $?"
tail -4 gpurun_out/r2f_bench_2gpu_phases.txt
timeout 1200 $TR bench.py --gpus 2 --steps 2 --warmup 1 --only-large > gpurun_out/r2f_large_2gpu.json 2> gpurun_out/r2f_large_2gpu.err; echo "large rc $?"
tail -3 gpurun_out/r2f_large_2gpu.err
python - <<'PY'
import json
b=json.load(open('gpurun_out/r2f_bench_2gpu.json'))
print({k:b[k] for k in ('value','ms_per_step','digest_ok','digest_mismatch')}, b['e2e']['ms_per_step'], b['configs1']['ms_per_step'], b['configs1']['digest_ok'])
l=json.load(open('gpurun_out/r2f_large_2gpu.json'))['large']
print({k:l[k] for k in ('value','ms_per_step','count_passes','batches_per_rank','phases_s_rank0','generation_s')}, l['output_digest'])
PY
