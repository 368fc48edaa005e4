#!/usr/bin/env bash
# range coder recode path (byte-exactness + timing), parity subset, and where a CLI run spends its time
mkdir -p gpurun_out
./tools/bin/rc_bench 92 50000 2>&1 | tail -3 | tee gpurun_out/r2l_rc_bench.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
python - <<'PY'
import os, subprocess, sys, time
sys.path.insert(0, '.')
from leon_b200 import workloads as W
b, o, q = W.generate("configs[2]")
os.makedirs("/tmp/f", exist_ok=True)
W.write_fixed("/tmp/f/w.fastq", b, 100, q)
cli = "leon_b200/bin/leon"
for conn in ("8", "32"):
    for rep in range(2):
        env = dict(os.environ, CUDA_DEVICE_MAX_CONNECTIONS=conn, LEON_B200_TRACE="1")
        t0 = time.perf_counter()
        r = subprocess.run([cli, "-file", "/tmp/f/w.fastq", "-c", "-noheader", "-out", "/tmp/f/o.leon"], env=env, capture_output=True, text=True)
        wall = time.perf_counter() - t0
        stages = [l.strip() for l in r.stdout.split("\n") if "Stages" in l]
        print("connections", conn, "rep", rep, "wall %.2f s" % wall, stages, [l for l in r.stderr.split("\n") if l][:3], flush=True)
import leon_b200
for rep in range(3):
    t0 = time.perf_counter()
    st = leon_b200.compress_file("/tmp/f/w.fastq", "/tmp/f/o2.leon", noheader=True)
    print("in-process %.3f s" % (time.perf_counter() - t0), {k: round(st[k], 1) for k in ("parse_ms", "compress_ms", "write_ms", "total_ms")}, "gpu total %.1f" % st["gpu"]["total_ms"], flush=True)
PY
