N=$1
timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err; tail -c 300 gpurun_out/r02_bench_${N}gpu.err | grep -v "^\*\|OMP_NUM" 
python - $N <<'PY'
import json,sys
d=json.loads(open('gpurun_out/r02_bench_%sgpu.json'%sys.argv[1]).read().strip().splitlines()[-1])
print('N',sys.argv[1],'value',d['value'],'e2e',d['e2e']['value'])
s=d['soup_mixed_10m']; print('soup',s['value'],s['generations_per_s_by_exchange'],s['host_enqueue_ms_per_generation'],s.get('sharded_equals_single_owner_bitwise'))
print('self_train',d['self_train']['value'],'kfv',d['kfv_sweep']['value'])
PY
