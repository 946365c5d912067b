O=gpurun_out; T=r2g
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py -m gpu -x -q --timeout 300 -k "ntt or witness_map or prove" > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -6 $O/${T}_pytest.log
timeout 300 python tools/bench_ntt.py > $O/${T}_ntt_tma.json 2> $O/${T}_ntt_tma.err; echo "tma $?"; cat $O/${T}_ntt_tma.json; tail -3 $O/${T}_ntt_tma.err
CTG_NTT_LEGACY=1 timeout 300 python tools/bench_ntt.py > $O/${T}_ntt_legacy.json 2> $O/${T}_ntt_legacy.err; echo "legacy $?"; cat $O/${T}_ntt_legacy.json
