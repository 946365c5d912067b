O=gpurun_out; T=r2z
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_circuits.py tests/test_host_cpp.py tests/test_codec.py -m gpu -x -q --timeout 300 -k "prove or round_trip or chain" > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
for lg in 12 16 18 20; do for v in 0 24; do echo "WM_FIRST_MAX_LOG=$v prove 2^$lg"; CTG_WM_FIRST_MAX_LOG=$v REPS=9 timeout 300 python tools/run_once.py prove $lg 0 2>&1 | grep total_ms | sed "s/.*'total_ms': \([0-9.]*\).*/\1/" | sort -n | head -6 | tr '\n' ' '; echo; done; done
