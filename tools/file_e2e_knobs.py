"""Wall clock of `leon -file X -c` on the configs[2] file with the scheduling knobs on and off (a fresh process and context per
run: first use of every buffer).  Usage: python tools/file_e2e_knobs.py > profiles/<tag>_file_e2e_knobs.txt"""
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from leon_b200 import workloads as W  # noqa: E402

name = "configs[2]"
tmp = tempfile.mkdtemp(prefix="leonknobs_")
path, n = bench.write_workload_file(name, tmp, 0)
cli = os.path.join(ROOT, "leon_b200", "bin", "leon")
flags = W.CONFIGS[name]["ref_flags"]
base = {k: v for k, v in os.environ.items() if k != "CUDA_DEVICE_MAX_CONNECTIONS"}
variants = [("default", {}), ("quals early", {"LEON_B200_QUALS_LATE": "0"}), ("no exclusive SMs", {"LEON_B200_RC_EXCLUSIVE": "0"}),
            ("msd 256", {"LEON_B200_MSD_THREADS": "256"}), ("default again", {}), ("trace", {"LEON_B200_TRACE": "1"})]
for label, extra in variants:
    walls = []
    for _ in range(3):
        env = dict(base)
        env.update(extra)
        t0 = time.perf_counter()
        r = subprocess.run([cli, "-file", path, "-c", "-kmer-size", "31", "-verbose", "1" if "TRACE" in str(extra) else "0", "-out", path + ".leon"] + flags,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
        walls.append(time.perf_counter() - t0)
        if r.returncode != 0:
            print(label, "FAILED", r.stdout[-300:])
            break
    print("%-18s" % label, " ".join("%.3f" % w for w in walls), "s")
    if "TRACE" in str(extra):
        print(r.stdout[-3000:])
