O=gpurun_out; T=r2e
mkdir -p $O
(time timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 300 -k "witness_batch or 1024_independent" --durations=5) > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"
tail -15 $O/${T}_pytest.log
(time timeout 1500 python bench.py) > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"; tail -5 $O/${T}_bench.err; python - <<'PY'
import json
l=json.load(open('gpurun_out/r2e_bench.json'))
ex=l.pop('extra',{})
print(json.dumps(l)[:1800])
for k,v in ex.items(): print(k, json.dumps(v)[:900])
PY
