#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 600 python bench.py --only-large --large-shape configs4 --large-reads 2000000 --large-passes 3 --steps 1 > gpurun_out/r2n_c4_small.json 2> gpurun_out/r2n_c4_small.err; echo "rc $?"
tail -3 gpurun_out/r2n_c4_small.err
python - <<'PY'
import json
l=json.load(open('gpurun_out/r2n_c4_small.json'))['large']
print({k:l[k] for k in ('value','ms_per_step','count_passes','kmer_size','quality','reads','genome','peak_device_bytes_per_rank','phases_s_rank0','cutoff','solid','anchors')})
PY
timeout 600 python bench.py --only-large --large-shape configs4 --large-reads 4000000 --large-passes 6 --steps 1 > gpurun_out/r2n_c4_small2.json 2>/dev/null
python - <<'PY'
import json
l=json.load(open('gpurun_out/r2n_c4_small2.json'))['large']
print({k:l[k] for k in ('value','ms_per_step','count_passes','reads','peak_device_bytes_per_rank','phases_s_rank0')})
PY
