O=gpurun_out; T=r2p
mkdir -p $O
nvidia-smi -L > $O/${T}_smi.txt
(time timeout 600 python -c "import __graft_entry__ as g; g.smoke()") > $O/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/${T}_smoke.log
(time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 10 --warmup 3) > $O/${T}_bench_n8.json 2> $O/${T}_bench_n8.err; echo "bench n8 exit $?"; tail -5 $O/${T}_bench_n8.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r2p_bench_n8.json').read().strip().splitlines()[-1])
ex=l.pop('extra',{})
print(json.dumps({k:l[k] for k in ('value','ms_per_step','e2e','matches_closed_form','n_gpus')}))
for k,v in ex.items(): print(k, json.dumps(v)[:1500])
PY
timeout 600 python -m pytest tests/test_gpu_multi_device.py -m gpu -x -q --timeout 200 > $O/${T}_pytest_multi.log 2>&1; echo "pytest-multi exit $?"; tail -3 $O/${T}_pytest_multi.log
