#!/usr/bin/env bash
# the no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (it collects last)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2j_noalloc_alone.log 2>&1; echo "alone rc $?" >> gpurun_out/r2j_noalloc_alone.log
tail -6 gpurun_out/r2j_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2j_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2j_pytest_gpu.log
tail -6 gpurun_out/r2j_pytest_gpu.log
