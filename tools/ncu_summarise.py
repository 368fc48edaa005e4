#!/usr/bin/env python3
"""Condenses the two ncu outputs of a GPU visit into the small tracked files under profiles/.

    python tools/ncu_summarise.py <tag> [gpurun_out/launches.csv] [gpurun_out/full.ncu-rep]

  profiles/<tag>_launches.csv    per kernel: launches, total / mean / max gpu__time_duration (us), share of the listed time
  profiles/<tag>_ncu_full.csv    per captured launch: duration, dram bytes read/written, dram / L2 / L1 throughput % of
                                 peak, L2 and L1 hit rates, achieved occupancy, issue-slot utilisation, registers
  profiles/traffic.json          kernel -> dram__bytes_read.sum + dram__bytes_write.sum per launch (bench.py roofline.traffic)
"""
import csv
import io
import json
import os
import re
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def short(name):
    m = re.search(r"(k_[a-z0-9_]+)", name)
    s = m.group(1) if m else name[:40]
    if "u128" in name:
        s += "<u128>"
    return s


def launches(path, tag):
    rows = [ln for ln in open(path) if ln.startswith('"')]
    rd = csv.DictReader(io.StringIO("".join(rows)))
    agg = OrderedDict()
    for r in rd:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] == "ns":
            v /= 1000.0
        elif r["Metric Unit"] == "ms":
            v *= 1000.0
        a = agg.setdefault(short(r["Kernel Name"]), [0, 0.0, 0.0])
        a[0] += 1
        a[1] += v
        a[2] = max(a[2], v)
    tot = sum(a[1] for a in agg.values())
    out = os.path.join(ROOT, "profiles", tag + "_launches.csv")
    with open(out, "w") as f:
        f.write("kernel,launches,total_us,mean_us,max_us,share\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("%s,%d,%.1f,%.1f,%.1f,%.4f\n" % (k, a[0], a[1], a[1] / a[0], a[2], a[1] / tot))
    print("wrote", out, "(%d launches, %.1f ms listed)" % (sum(a[0] for a in agg.values()), tot / 1000))


WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram_read"),
    ("dram__bytes_write.sum", "dram_write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_pct"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1_pct"),
    ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
    ("l1tex__t_sector_hit_rate.pct", "l1_hit_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
    ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue_pct"),
    ("smsp__inst_executed.sum", "warp_insts"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
]
UNIT_TO_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
UNIT_TO_US = {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}


def full(path, tag):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {}
    for i, h in enumerate(hdr):
        col.setdefault(h, i)
    out = os.path.join(ROOT, "profiles", tag + "_ncu_full.csv")
    traffic = {}
    with open(out, "w") as f:
        f.write("kernel," + ",".join(n for _, n in WANT) + "\n")
        for r in data:
            vals = []
            k = short(r[col["Kernel Name"]])
            rec = {}
            for metric, name in WANT:
                i = col.get(metric)
                if i is None or r[i] in ("", "no data"):
                    vals.append("")
                    continue
                v = float(r[i].replace(",", ""))
                u = units[i]
                if name in ("dram_read", "dram_write"):
                    v *= UNIT_TO_BYTES.get(u, 1)
                    vals.append("%d" % v)
                elif name == "duration":
                    v *= UNIT_TO_US.get(u, 1)
                    vals.append("%.1f" % v)
                else:
                    vals.append("%.2f" % v if v != int(v) else "%d" % v)
                rec[name] = v
            f.write(k + "," + ",".join(vals) + "\n")
            if "dram_read" in rec and "dram_write" in rec:
                traffic.setdefault(k, []).append(rec["dram_read"] + rec["dram_write"])
    print("wrote", out)
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    old = json.load(open(tpath)) if os.path.exists(tpath) else {}
    for k, v in traffic.items():
        old[k] = max(v)     # the largest launch of that kernel (k_part_scatter: the k-mer pass, not the symbol-record pass)
    json.dump(old, open(tpath, "w"), indent=1, sort_keys=True)
    print("wrote", tpath, old)


if __name__ == "__main__":
    tag = sys.argv[1]
    lp = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "launches.csv")
    fp = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", "full.ncu-rep")
    if os.path.exists(lp):
        launches(lp, tag)
    if os.path.exists(fp):
        full(fp, tag)
