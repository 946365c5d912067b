O=gpurun_out; T=r2z8
mkdir -p $O
(time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3) > $O/${T}_bench_n2.json 2> $O/${T}_bench_n2.err; echo "bench n2 exit $?"; tail -4 $O/${T}_bench_n2.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r2z8_bench_n2.json').read().strip().splitlines()[-1])
ex=l.pop('extra',{})
print(json.dumps({k:l[k] for k in ('value','ms_per_step','matches_closed_form','n_gpus')}), l['e2e']['value'])
for k,v in ex.items(): print(k, json.dumps(v)[:600])
PY
