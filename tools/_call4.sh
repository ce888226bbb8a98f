export PROF_ONLY_WW=1
for v in train1 dev_mb4 dev_mb5 dev_mb6; do echo "== $v"; SRNN_LIB=$PWD/build/variants/$v.so python tools/prof_train.py 20 100 5 2>&1 | tail -1; done | tee gpurun_out/r02_train_ab2.txt
SRNN_LIB=$PWD/build/variants/dev_mb5.so python -m pytest tests/test_gpu_sgd_pin.py -x -q -m gpu 2>&1 | tail -2
