O=gpurun_out; T=r2n
mkdir -p $O
(time timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300) > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $O/${T}_pytest.log
(time timeout 1500 python bench.py) > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"; tail -5 $O/${T}_bench.err; python - <<'PY'
import json
l=json.load(open('gpurun_out/r2n_bench.json'))
ex=l.pop('extra',{})
print(json.dumps(l)[:600])
for k,v in ex.items():
    if k.startswith('batched') or k.startswith('g1_msm') or k.startswith('ntt_fr') or k.startswith('g2'): print(k, json.dumps(v)[:1000])
    elif k.startswith('groth16'): print(k, v.get('wall_ms_median'), v.get('cpu_port_prove_ms'), v.get('gpu_matches_cpu'))
PY
