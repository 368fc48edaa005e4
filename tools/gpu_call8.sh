#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tools/multigpu_check.py > gpurun_out/mg_check_$N.log 2>&1; echo "mgcheck rc=$?"; tail -6 gpurun_out/mg_check_$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --steps 3 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err; echo "bench$N rc=$?"; grep -E "phases|kernels" gpurun_out/bench_$N.err | cut -c1-700; cut -c1-260 gpurun_out/bench_$N.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --fastq --steps 3 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_fastq_$N.json 2> gpurun_out/bench_fastq_$N.err; echo "bench fastq $N rc=$?"; cut -c1-260 gpurun_out/bench_fastq_$N.json
