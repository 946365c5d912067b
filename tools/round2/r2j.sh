O=gpurun_out; T=r2j
mkdir -p $O
nvidia-smi -L > $O/${T}_smi.txt
(time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3) > $O/${T}_bench_n2.json 2> $O/${T}_bench_n2.err; echo "bench n2 exit $?"; tail -5 $O/${T}_bench_n2.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r2j_bench_n2.json').read().strip().splitlines()[-1])
ex=l.pop('extra',{})
print(json.dumps(l)[:1500])
for k,v in ex.items(): print(k, json.dumps(v)[:1200])
PY
timeout 600 python -m pytest tests/test_gpu_multi_device.py -m gpu -x -q --timeout 200 > $O/${T}_pytest_multi.log 2>&1; echo "pytest-multi exit $?"; tail -5 $O/${T}_pytest_multi.log
