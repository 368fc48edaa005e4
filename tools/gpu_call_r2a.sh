#!/usr/bin/env bash
# round 2, GPU visit a: parity suite, range coder micro-benchmark (v1 vs v2), bench both arms
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_gpu.txt
lscpu | head -20 > gpurun_out/r2a_cpu.txt; free -g >> gpurun_out/r2a_cpu.txt
./tools/bin/rc_bench 92 50000 > gpurun_out/r2a_rc_bench.txt 2>&1
./tools/bin/lat_bench > gpurun_out/r2a_lat_bench.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2a_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --phases > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench_phases.txt; echo "bench rc $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_bench_reference.json 2> gpurun_out/r2a_bench_reference.err; echo "ref rc $?"
tail -3 gpurun_out/r2a_pytest_gpu.log; cat gpurun_out/r2a_rc_bench.txt; head -c 1500 gpurun_out/r2a_bench.json
