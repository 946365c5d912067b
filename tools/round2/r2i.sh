O=gpurun_out; T=r2i
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300 > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $O/${T}_pytest.log
timeout 300 python tools/bench_ntt.py 16 20 22 24 > $O/${T}_ntt_tma.json 2> $O/${T}_ntt_tma.err; echo "tma $?"; cat $O/${T}_ntt_tma.json; tail -3 $O/${T}_ntt_tma.err
CTG_NTT_LEGACY=1 timeout 300 python tools/bench_ntt.py 20 22 24 > $O/${T}_ntt_legacy.json 2> $O/${T}_ntt_legacy.err; echo "legacy $?"; cat $O/${T}_ntt_legacy.json
timeout 300 python tools/bench_poseidon.py > $O/${T}_poseidon.json 2>&1; tail -5 $O/${T}_poseidon.json
