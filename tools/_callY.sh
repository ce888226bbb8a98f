export PROF_ONLY_WW=1
for v in "" build/variants/dev_packT.so; do
  if [ -n "$v" ]; then export SRNN_LIB=$PWD/$v; fi
  echo "== lib ${v:-tree}"
  python tools/prof_train.py 20 100 5
  python tools/prof_soup.py 10000000 20 0
  python tools/prof_soup.py 10000000 20 1
  PROF_LEARN=-1 python tools/prof_soup.py 100000 400 1
  PROF_LEARN=-1 python tools/prof_soup.py 100000 100 30
  python tools/prof_soup.py 10000000 10 0 0 exact
done
SRNN_LIB=$PWD/build/variants/dev_packT.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sgd_pin.py -q -m gpu -k "train or sgd or learn" 2>&1 | tail -3
