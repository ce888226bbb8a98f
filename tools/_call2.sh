set -x
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_pytest_gpu_1.log 2>&1; tail -5 gpurun_out/r02_pytest_gpu_1.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_1.json 2> gpurun_out/r02_bench_1.err; tail -c 600 gpurun_out/r02_bench_1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_1.json').read().strip().splitlines()[-1])
for k in ('value','e2e','roofline'): print(k, d.get(k))
print('self_train', d['self_train']['value'], d['self_train']['roofline'])
print('kfv', d['kfv_sweep']['value']); print('mixed', d['soup_mixed_10m']['value'], d['soup_mixed_10m']['host_enqueue_ms_per_generation'])
for k in ('soup','soup_exact_order','soup_10m','soup_10m_exact_order','soup_train30'): print(k, d[k]['value'])
print('facade', d.get('facade'))
print('cpu', d.get('cpu_baseline'))
PY
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_ref_1.json 2>&1; cut -c1-400 gpurun_out/r02_bench_ref_1.json
