# same-box A/B of libsrnn_b200.so variants placed under build/variants/ (SRNN_LIB selects the library)
for v in old ppt2_mb6; do
  export SRNN_LIB=$PWD/build/variants/$v.so
  echo "=== $v"
  for r in "-1 -1" "0.1 -1" "-1 0.1" "0.1 0.1" "0.3 -1" "-1 0.3"; do
    set -- $r
    PROF_ATTACK=$1 PROF_LEARN=$2 timeout 120 python tools/prof_soup.py 10000000 20 0 2>&1 | tail -1 | cut -c1-100
  done
done
