set -x
for v in ppt2_mb6 ppt3_mb4 ppt3_mb5; do
  export SRNN_LIB=$PWD/build/variants/$v.so
  echo "=== $v"
  timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "soup" 2>&1 | tail -2
  for args in "10000000 20 0" "10000000 10 1" "100000 200 0" "1000000 50 0" "1000000 20 1"; do
    timeout 120 python tools/prof_soup.py $args 2>&1 | tail -1
  done
done
