#!/usr/bin/env bash
# 8 GPUs: byte parity against the oracle (8 cases incl. streaming, k = 63 lossless), the driver's bench command at N = 8
# (configs[2] + configs[1] + configs[3] streamed, digests), and the configs[4]-shaped sample (k = 63, lossless, streamed in
# batches of 1 M reads per rank) at 30 M and at 15 M reads of the same 150 Mbp genome: peak device memory must not follow R
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/r2o_gpus.txt; nproc >> gpurun_out/r2o_gpus.txt; free -g | head -2 >> gpurun_out/r2o_gpus.txt
timeout 600 $TR tools/multigpu_check.py > gpurun_out/r2o_multigpu_check_8gpu.log 2>&1; echo "check rc $?"
grep "^case\|ALL OK\|FAILED\|MISMATCH" gpurun_out/r2o_multigpu_check_8gpu.log
timeout 800 $TR bench.py --gpus 8 --steps 5 --warmup 3 --phases > gpurun_out/r2o_bench_8gpu.json 2> gpurun_out/r2o_bench_8gpu_phases.txt; echo "bench rc $?"
grep "^phases" gpurun_out/r2o_bench_8gpu_phases.txt | head -1
timeout 600 $TR bench.py --gpus 8 --only-large --large-shape configs4 --large-batch-chunks 4 --steps 1 > gpurun_out/r2o_c4_30M_8gpu.json 2> gpurun_out/r2o_c4_30M.err; echo "c4 30M rc $?"
timeout 600 $TR bench.py --gpus 8 --only-large --large-shape configs4 --large-batch-chunks 4 --large-reads 15000000 --large-genome 150000000 --steps 1 > gpurun_out/r2o_c4_15M_8gpu.json 2> gpurun_out/r2o_c4_15M.err; echo "c4 15M rc $?"
python - <<'PY'
import json
def last(p):
    return json.loads([l for l in open(p) if l.startswith('{')][-1])
b=last('gpurun_out/r2o_bench_8gpu.json')
print({k:b[k] for k in ('value','ms_per_step','digest_ok','digest_mismatch')}, 'e2e', b['e2e']['ms_per_step'], 'c1', b['configs1']['ms_per_step'], b['configs1']['digest_ok'])
l=b['large']; print('c3', {k:l[k] for k in ('value','ms_per_step','count_passes','batches_per_rank','phases_s_rank0','peak_device_bytes_per_rank','digest_ok')}, l['output_digest'])
for f in ('r2o_c4_30M_8gpu','r2o_c4_15M_8gpu'):
    l=last('gpurun_out/%s.json'%f)['large']
    print(f, {k:l[k] for k in ('value','ms_per_step','count_passes','batches_per_rank','reads','genome','peak_device_bytes_per_rank','phases_s_rank0','cutoff','solid','anchors','bloom_bytes')})
PY
