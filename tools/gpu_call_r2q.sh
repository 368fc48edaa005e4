#!/usr/bin/env bash
mkdir -p gpurun_out
for i in 1 2; do ./tools/bin/rc_bench --dict 300000 31; ./tools/bin/rc_bench_v3 --dict 300000 31 | sed 's/^/x86-64-v3 /'; done | tee gpurun_out/r2q_dict_coder_host.txt
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2q_pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2q_pytest_gpu.log
tail -4 gpurun_out/r2q_pytest_gpu.log
timeout 1500 python bench.py --steps 10 --warmup 3 --phases > gpurun_out/r2q_bench.json 2> gpurun_out/r2q_bench_phases.txt; echo "bench rc $?"
python - <<'PY'
import json
b=json.load(open('gpurun_out/r2q_bench.json'))
print({k:b[k] for k in ('value','ms_per_step','digest_ok')}, 'e2e', b['e2e']['ms_per_step'], 'c1', b['configs1']['ms_per_step'], b['configs1']['digest_ok'])
print(b['cpu_baseline']['value'], b['file_e2e'])
l=b['large']; print({k:l[k] for k in ('value','ms_per_step','count_passes','batches_per_rank','phases_s_rank0','peak_device_bytes_per_rank','digest_ok')})
PY
# device memory of the streaming path against the number of reads: same genome, same batches (500 k reads), passes scaled with R
for cfg in "2000000 2" "4000000 4" "8000000 8"; do set -- $cfg
  timeout 600 python bench.py --only-large --large-shape configs4 --large-reads $1 --large-genome 20000000 --large-batch-chunks 2 --large-passes $2 --steps 1 2>/dev/null | python -c "
import json,sys
l=json.loads([x for x in sys.stdin if x.startswith('{')][-1])['large']
print('reads %d passes %d batches %d peak device bytes %d  %.0f ms' % (l['reads'], l['count_passes'], l['batches_per_rank'], l['peak_device_bytes_per_rank'], l['ms_per_step']))"
done | tee gpurun_out/r2q_stream_memory.txt
