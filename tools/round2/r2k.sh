O=gpurun_out; T=r2k
mkdir -p $O
timeout 300 python tools/bench_ntt.py 20 22 24 > $O/${T}_ntt_r8.json 2> $O/${T}_ntt_r8.err; echo "r8 $?"; cat $O/${T}_ntt_r8.json
CTG_NTT_RADIX4=1 timeout 300 python tools/bench_ntt.py 20 22 24 > $O/${T}_ntt_r4.json 2> $O/${T}_ntt_r4.err; echo "r4 $?"; cat $O/${T}_ntt_r4.json; tail -2 $O/${T}_ntt_r4.err
CTG_NTT_RADIX4=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py -m gpu -x -q --timeout 300 -k "ntt or witness_map" > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
F="ncu --set full --clock-control none --import-source on -f"
REPS=1 timeout 600 $F -k regex:"ntt_pass" -s 0 -c 3 -o $O/${T}_ntt_r8 python tools/run_once.py ntt 22 > $O/${T}_ncu_r8.log 2>&1; echo "ncu $?"
python tools/ncu_summary.py $O/${T}_ntt_r8.ncu-rep > $O/${T}_ntt_r8_summary.txt 2>&1
CTG_NTT_RADIX4=1 REPS=1 timeout 600 $F -k regex:"ntt_pass" -s 0 -c 3 -o $O/${T}_ntt_r4 python tools/run_once.py ntt 22 > $O/${T}_ncu_r4.log 2>&1; echo "ncu $?"
python tools/ncu_summary.py $O/${T}_ntt_r4.ncu-rep > $O/${T}_ntt_r4_summary.txt 2>&1
rm -f $O/${T}_ntt_r4.ncu-rep
ls -la $O/${T}_ntt_r8.ncu-rep
