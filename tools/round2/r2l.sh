O=gpurun_out; T=r2l
mkdir -p $O
timeout 300 python tools/bench_ntt.py 20 22 24 > $O/${T}_ntt_r8.json 2> $O/${T}_ntt_r8.err; echo "r8 $?"; cat $O/${T}_ntt_r8.json
CTG_NTT_RADIX4=1 timeout 300 python tools/bench_ntt.py 20 22 24 > $O/${T}_ntt_r4.json 2> $O/${T}_ntt_r4.err; echo "r4 $?"; cat $O/${T}_ntt_r4.json
timeout 300 python tools/bench_poseidon.py > $O/${T}_poseidon.json 2>&1; tail -2 $O/${T}_poseidon.json | cut -c1-900
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 -k "ntt or witness_map or field or poseidon or elgamal or te_mul or prove_batch" > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
L="ncu --metrics gpu__time_duration.sum --clock-control none --csv"
REPS=1 timeout 600 $L -k regex:"^(?!.*field_op)" -c 1200 --log-file $O/${T}_batch_launches.csv python tools/run_once.py batch 256 > $O/${T}_batch_ncu.log 2>&1; echo "batch launches $?"
python tools/launch_summary.py $O/${T}_batch_launches.csv > $O/${T}_batch_launches_summary.txt; head -40 $O/${T}_batch_launches_summary.txt
