O=gpurun_out; T=r2t
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py tests/test_gpu_multi_device.py -m gpu -x -q --timeout 300 -k "g2 or prove or msm or fq2 or field" > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
for v in 1 0; do for lg in 16 20; do echo "CTG_G2_PAIR=$v msm2 2^$lg"; CTG_G2_PAIR=$v REPS=3 timeout 300 python tools/run_once.py msm2 $lg 0 2>&1 | grep msm_sort | tail -1 | sed "s/.*'msm_sort_ms'/sort/"; done; done
for v in 1 0; do echo "CTG_G2_PAIR=$v prove 2^16"; CTG_G2_PAIR=$v REPS=8 timeout 300 python tools/run_once.py prove 16 0 2>&1 | grep total_ms | sed "s/.*'total_ms': \([0-9.]*\).*/\1/" | sort -n | tr '\n' ' '; echo; done
for v in 1 0; do echo "CTG_G2_PAIR=$v prove 2^20"; CTG_G2_PAIR=$v REPS=8 timeout 300 python tools/run_once.py prove 20 0 2>&1 | grep total_ms | sed "s/.*'total_ms': \([0-9.]*\).*/\1/" | sort -n | tr '\n' ' '; echo; done
