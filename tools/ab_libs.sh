# same-box A/B of libsrnn_b200.so variants under build/variants/ plus the in-tree build ("tree"); SRNN_LIB selects
for v in "$@"; do
  if [ "$v" = tree ]; then unset SRNN_LIB; else export SRNN_LIB=$PWD/build/variants/$v.so; fi
  echo "=== $v"
  for args in "10000000 20 0" "10000000 10 1" "1000000 50 0" "1000000 50 1" "100000 200 0" "100000 200 1"; do
    timeout 120 python tools/prof_soup.py $args 2>&1 | tail -1 | cut -c1-110
  done
  PROF_ATTACK=-1 PROF_LEARN=0.1 timeout 120 python tools/prof_soup.py 10000000 20 0 2>&1 | tail -1 | cut -c1-110
  PROF_ATTACK=0.1 PROF_LEARN=-1 timeout 120 python tools/prof_soup.py 10000000 20 0 2>&1 | tail -1 | cut -c1-110
  PROF_ATTACK=-1 PROF_LEARN=-1 timeout 120 python tools/prof_soup.py 10000000 20 0 2>&1 | tail -1 | cut -c1-110
done
