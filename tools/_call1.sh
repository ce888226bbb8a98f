set -x
mkdir -p gpurun_out
nvidia-smi -L
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r02_pytest_gpu_0.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu_0.log
python tools/facade_bench.py > gpurun_out/r02_facade_0.json 2> gpurun_out/r02_facade_0.err; cat gpurun_out/r02_facade_0.json; tail -3 gpurun_out/r02_facade_0.err
python tools/run_ref_setups.py --timeout 600 --out gpurun_out/r02_ref_setups_0.json > gpurun_out/r02_ref_setups_0.log 2>&1; python - <<'PY'
import json
for r in json.load(open('gpurun_out/r02_ref_setups_0.json')):
    print(r['script'], r['rc'], r['seconds'])
    if r['rc']: print(r['tail'][-800:])
PY
