O=gpurun_out; T=r2s
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py -m gpu -x -q --timeout 300 -k "ntt or witness_map or prove" > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
timeout 300 python tools/bench_ntt.py 13 16 19 20 22 24 > $O/${T}_ntt_t10.json 2> $O/${T}_ntt_t10.err; echo "t10 $?"; cat $O/${T}_ntt_t10.json
CTG_NTT_TILE=11 timeout 300 python tools/bench_ntt.py 13 16 19 20 22 24 > $O/${T}_ntt_t11.json 2> $O/${T}_ntt_t11.err; echo "t11 $?"; cat $O/${T}_ntt_t11.json
for v in "CTG_NTT_TILE=10" "CTG_NTT_TILE=11" "CTG_NTT_LEGACY=1" "CTG_NTT_TILE=10"; do echo "$v 2^22"; env $v REPS=8 timeout 300 python tools/run_once.py prove 22 0 2>&1 | grep total_ms | sed "s/.*'total_ms': \([0-9.]*\).*/\1/" | sort -n | tr '\n' ' '; echo; done
for v in "CTG_NTT_TILE=10" "CTG_NTT_LEGACY=1"; do echo "$v 2^20"; env $v REPS=8 timeout 300 python tools/run_once.py prove 20 0 2>&1 | grep total_ms | sed "s/.*'total_ms': \([0-9.]*\).*/\1/" | sort -n | tr '\n' ' '; echo; done
