export SRNN_LIB=$PWD/build/variants/dev_skip.so
python bench.py --no-legs 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('headline %.4e  e2e %.3e  frac %.3f' % (d['value'], d['e2e']['value'], d['roofline']['frac']))"
timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "self_apply or golden or nan_inf" 2>&1 | tail -5
