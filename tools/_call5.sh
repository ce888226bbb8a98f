export PROF_ONLY_WW=1
for v in train1 dev2 dev3; do echo "== $v"; SRNN_LIB=$PWD/build/variants/$v.so python tools/prof_train.py 20 100 5 2>&1 | tail -1; done | tee gpurun_out/r02_train_ab3.txt
SRNN_LIB=$PWD/build/variants/dev3.so python -m pytest tests/test_gpu_sgd_pin.py -x -q -m gpu 2>&1 | tail -2
SRNN_LIB=$PWD/build/variants/dev3.so ncu --set full --clock-control none --import-source on -k regex:train_kernel -c 2 -o gpurun_out/r02_train_ww_dev3 python tools/prof_train.py 20 10 1 > gpurun_out/r02_train_ncu3.log 2>&1
ncu -i gpurun_out/r02_train_ww_dev3.ncu-rep --page raw --csv > gpurun_out/r02_train_ww_dev3_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r02_train_ww_dev3_raw.csv
