O=gpurun_out; T=r2z6
mkdir -p $O
timeout 600 python -m pytest tests/test_circuits.py -m gpu -x -q --timeout 300 > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
(time timeout 1500 python bench.py) > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"; tail -3 $O/${T}_bench.err; python - <<'PY'
import json
l=json.load(open('gpurun_out/r2z6_bench.json')); ex=l['extra']
print(l['value'], l['e2e']['value'], l['roofline']['frac'], l['matches_closed_form'])
for k,v in ex.items():
    if k.startswith('batched'): print(k, json.dumps(v)[:420])
PY
