#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
./tools/bin/rc_bench 92 50000 > gpurun_out/rc_bench.log 2>&1; echo "rc_bench rc=$?"; cat gpurun_out/rc_bench.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -4 gpurun_out/bench.err
