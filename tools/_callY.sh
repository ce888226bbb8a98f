timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -8
timeout 600 python bench.py > gpurun_out/bench_n3.json 2> gpurun_out/bench_n3.err; echo bench rc=$?
tail -c 600 gpurun_out/bench_n3.err
