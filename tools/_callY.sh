python tools/prof_mixed.py
python tools/prof_soup.py 10000000 20 0
python tools/prof_soup.py 10000000 20 1
PROF_KIND=1 python tools/prof_soup.py 3333328 20 0
PROF_KIND=2 python tools/prof_soup.py 3333328 20 0
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_sharded_soup.py tests/test_gpu_facade.py -q -m gpu -k "soup or evolve" 2>&1 | tail -5
