# same-box A/B of libsrnn_b200.so variants under build/variants/ plus the in-tree build ("tree"); SRNN_LIB selects
for v in "$@"; do
  if [ "$v" = tree ]; then unset SRNN_LIB; else export SRNN_LIB=$PWD/build/variants/$v.so; fi
  echo "=== $v"
  python tools/prof_apply.py 24 2>&1 | grep -v "^$" | tail -9
  for args in "10000000 20 0" "10000000 10 1" "1000000 50 0" "100000 200 0"; do
    timeout 120 python tools/prof_soup.py $args 2>&1 | tail -1 | cut -c1-110
  done
  timeout 120 python tools/prof_soup.py 100000 100 1 0 exact 2>&1 | tail -1 | cut -c1-110
  timeout 120 python tools/prof_soup.py 10000000 5 0 0 exact 2>&1 | tail -1 | cut -c1-110
done
