O=gpurun_out; T=r2u
mkdir -p $O
for c in 16 17 18 19; do echo "CHUNK_LOG=$c"; CTG_REDUCE_CHUNK_LOG=$c REPS=3 timeout 300 python tools/run_once.py msm2 20 0 2>&1 | grep msm_sort | tail -1 | sed "s/.*'msm_sort_ms'/G2 2^20 sort/"; CTG_REDUCE_CHUNK_LOG=$c REPS=3 timeout 300 python tools/run_once.py msm 22 0 2>&1 | grep msm_sort | tail -1 | sed "s/.*'msm_sort_ms'/G1 2^22 sort/"; done
CTG_REDUCE_CHUNK_LOG=19 timeout 600 python -m pytest tests/test_gpu_baseline_sizes.py tests/test_gpu_parity.py -m gpu -x -q --timeout 300 -k "msm" > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
