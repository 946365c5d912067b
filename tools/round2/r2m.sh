O=gpurun_out; T=r2m
mkdir -p $O
timeout 600 python -m pytest tests/test_circuits.py -m gpu -x -q --timeout 300 > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $O/${T}_pytest.log
for c in 0 11 12 13 14 15; do echo "CBITS=$c"; CBITS=$c REPS=3 timeout 300 python tools/run_once.py batch 1020 2>&1 | grep prove_batch | tail -2; done
