O=gpurun_out; T=r2x
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py -m gpu -x -q --timeout 300 -k "msm or prove" > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
for v in 1 0; do for lg in 16 22; do echo "CTG_G1_PAIR=$v msm 2^$lg"; CTG_G1_PAIR=$v REPS=3 timeout 300 python tools/run_once.py msm $lg 0 2>&1 | grep msm_sort | tail -1 | sed "s/.*'msm_sort_ms'/sort/"; done; done
for v in 1 0; do echo "CTG_G1_PAIR=$v prove 2^16"; CTG_G1_PAIR=$v REPS=8 timeout 300 python tools/run_once.py prove 16 0 2>&1 | grep total_ms | sed "s/.*'total_ms': \([0-9.]*\).*/\1/" | sort -n | tr '\n' ' '; echo; done
for v in 1 0; do echo "CTG_G1_PAIR=$v batch 1020"; CTG_G1_PAIR=$v REPS=3 timeout 300 python tools/run_once.py batch 1020 2>&1 | grep prove_batch | tail -2; done
