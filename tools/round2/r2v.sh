O=gpurun_out; T=r2v
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${T}_smi.txt 2>&1; nproc >> $O/${T}_smi.txt
(time timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300) > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
(time timeout 1500 python bench.py) > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"; tail -3 $O/${T}_bench.err
(time timeout 900 python bench.py --impl reference --steps 2 --warmup 1) > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref exit $?"; head -c 900 $O/${T}_bench_ref.json; echo
L="ncu --metrics gpu__time_duration.sum --clock-control none --csv"
F="ncu --set full --clock-control none --import-source on -f"
timeout 600 $L -c 400 --log-file $O/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > $O/${T}_ncu_launch.log 2>&1; echo "ncu launches $?"
python tools/launch_summary.py $O/${T}_launches.csv > $O/${T}_launches_summary.txt
REPS=1 timeout 600 $F -k regex:"ntt_pass" -s 0 -c 3 -o $O/${T}_ntt python tools/run_once.py ntt 22 > $O/${T}_ncu_ntt.log 2>&1; echo "ncu ntt $?"
python tools/ncu_summary.py $O/${T}_ntt.ncu-rep > $O/${T}_ntt_summary.txt 2>&1; rm -f $O/${T}_ntt.ncu-rep
REPS=1 timeout 600 $F -k regex:"pair_kernel" -s 0 -c 8 -o $O/${T}_pair python tools/run_once.py msm2 20 0 > $O/${T}_ncu_pair.log 2>&1; echo "ncu pair $?"
python tools/ncu_summary.py $O/${T}_pair.ncu-rep > $O/${T}_pair_summary.txt 2>&1; rm -f $O/${T}_pair.ncu-rep
REPS=1 timeout 600 $L -c 200 --log-file $O/${T}_g2_launches.csv python tools/run_once.py msm2 20 0 > /dev/null 2>&1; python tools/launch_summary.py $O/${T}_g2_launches.csv > $O/${T}_g2_launches_summary.txt; rm -f $O/${T}_g2_launches.csv
du -sh $O
