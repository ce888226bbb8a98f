set -x
mkdir -p gpurun_out
python tools/prof_train.py 20 100 5 2>&1 | tee gpurun_out/r02_train_ab_tree.txt
SRNN_LIB=$PWD/build/variants/train1.so python tools/prof_train.py 20 100 5 2>&1 | tee gpurun_out/r02_train_ab_train1.txt
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sgd_pin.py -x -q -m gpu 2>&1 | tail -3
ncu --set full --clock-control none --import-source on -k regex:train_kernel -c 3 -o gpurun_out/r02_train_ww python tools/prof_train.py 20 10 1 > gpurun_out/r02_train_ncu.log 2>&1
ncu -i gpurun_out/r02_train_ww.ncu-rep --page raw --csv > gpurun_out/r02_train_ww_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r02_train_ww_raw.csv
ls -la gpurun_out/*.ncu-rep
