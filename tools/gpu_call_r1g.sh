#!/usr/bin/env bash
# GPU-box visit for the tree with the GPU decoder: decoder parity first, then the whole GPU suite, smoke, both bench
# arms, the decoder benchmark, launch list of the bench command and an ncu --set full capture of k_decode_blocks on a
# reduced case (92 blocks of 5000 reads).  Every ncu run comes after the same command has exited 0 without ncu.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py -m gpu -x -q > gpurun_out/pytest_gpu_decode.log 2>&1; echo "pytest decode rc=$?"; tail -2 gpurun_out/pytest_gpu_decode.log
timeout 600 python tools/decode_bench.py --out gpurun_out/decode_bench.json > gpurun_out/decode_bench.log 2>&1; echo "decode bench rc=$?"; cat gpurun_out/decode_bench.json
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 5 --warmup 3 --phases > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python bench.py --fastq --steps 5 --warmup 3 --phases --no-cpu-baseline > gpurun_out/bench_fastq.json 2> gpurun_out/bench_fastq.err; echo "bench fastq rc=$?"
if [ "${WITH_NCU:-1}" = 1 ]; then
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:k_decode_blocks' -c 1 -f -o gpurun_out/full_decode \
    python tools/decode_bench.py --reads 460000 --rpb 5000 > gpurun_out/ncu_full_decode.log 2>&1; echo "ncu full decode rc=$?"
fi
cut -c1-300 gpurun_out/bench.json; cut -c1-300 gpurun_out/bench_fastq.json
