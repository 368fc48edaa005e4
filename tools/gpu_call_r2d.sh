#!/usr/bin/env bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "streaming or pipelined or golden or seeded or edge" > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2d_pytest.log
tail -30 gpurun_out/r2d_pytest.log
