#!/usr/bin/env bash
# r1i: k_walk with 64-bit probe loads (7 LDG.64 + 7 mostly predicated-off LDG per step).  GPU suite, smoke, bench lines,
# configs[3]-scale run, launch list, ncu --set full of k_walk on the bench command.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 5 --warmup 3 --phases > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python bench.py --fastq --steps 5 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_fastq.json 2> gpurun_out/bench_fastq.err; echo "bench fastq rc=$?"
grep -h -o "ms_per_step\": [0-9.]*" gpurun_out/bench.json gpurun_out/bench_fastq.json
grep -h -o "'k_walk': ([0-9.]*\|'k_quals': ([0-9.]*" gpurun_out/bench.err gpurun_out/bench_fastq.err
timeout 600 python tools/run_scale.py --genome 100000000 --reads 33333333 --len 150 > gpurun_out/c4_run.log 2>&1; echo "c4 rc=$?"
grep -h -o '"k_walk": [0-9.]*\|digest.*\|"wall_s": [0-9.]*' gpurun_out/c4_run.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:k_walk|k_range_encode$' -c 2 -f -o gpurun_out/full_walk \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_full_walk.log 2>&1; echo "ncu full walk rc=$?"
