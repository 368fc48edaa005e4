#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench1 rc=$?"; tail -3 gpurun_out/bench1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench2 rc=$?"; tail -5 gpurun_out/bench2.err; cut -c1-300 gpurun_out/bench2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 tools/multigpu_check.py > gpurun_out/mg_check.log 2>&1; echo "mgcheck rc=$?"; tail -6 gpurun_out/mg_check.log
