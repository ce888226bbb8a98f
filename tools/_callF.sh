export SRNN_LIB=$PWD/build/variants/dev_defer.so
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "soup_sync_matches or full_size or chain_peers" 2>&1 | tail -4
for rates in "0.1 0.1" "-1 0.1"; do set -- $rates; PROF_KIND=2 PROF_ATTACK=$1 PROF_LEARN=$2 timeout 120 python tools/prof_soup.py 3333328 20 0 2>&1 | tail -1 | cut -c1-100; done
