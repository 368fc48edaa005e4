#!/usr/bin/env bash
# 2 GPUs on the final tree: byte parity against the oracle (8 cases incl. streaming), the driver's bench command at N = 2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 500 $TR tools/multigpu_check.py > gpurun_out/r2z_multigpu_check_2gpu.log 2>&1; echo "check rc $?"
grep "^case\|ALL OK\|FAILED\|MISMATCH" gpurun_out/r2z_multigpu_check_2gpu.log
timeout 700 $TR bench.py --gpus 2 --steps 5 --warmup 3 --phases > gpurun_out/r2z_bench_2gpu.json 2> gpurun_out/r2z_bench_2gpu_phases.txt; echo "bench rc $?"
python - <<'PY'
import json
def last(p):
    return json.loads([l for l in open(p) if l.startswith('{')][-1])
b=last('gpurun_out/r2z_bench_2gpu.json')
print({k:b[k] for k in ('value','ms_per_step','digest_ok','digest_mismatch')}, 'e2e', b['e2e']['ms_per_step'], 'c1', b['configs1']['ms_per_step'], b['configs1']['digest_ok'])
l=b['large']; print('c3', {k:l[k] for k in ('value','ms_per_step','count_passes','batches_per_rank','peak_device_bytes_per_rank','digest_ok','digest_mismatch')}); p=l['phases_s_rank0']; print({k:p[k] for k in ('count_s','bloom_s','encode_s','dict_host_s','dict_wait_s')})
PY
