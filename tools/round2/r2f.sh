O=gpurun_out; T=r2f
mkdir -p $O
L="ncu --metrics gpu__time_duration.sum --clock-control none --csv"
F="ncu --set full --clock-control none --import-source on -f"
sumrep() { python tools/ncu_summary.py $O/$1.ncu-rep > $O/$1_summary.txt 2>&1; rm -f $O/$1.ncu-rep; }
REPS=1 timeout 600 $L -c 3000 --log-file $O/${T}_batch_launches.csv python tools/run_once.py batch 256 > $O/${T}_batch_ncu.log 2>&1; echo "batch launches $?"
python tools/launch_summary.py $O/${T}_batch_launches.csv > $O/${T}_batch_launches_summary.txt; rm $O/${T}_batch_launches.csv
REPS=1 timeout 600 $L -c 600 --log-file $O/${T}_g2_launches.csv python tools/run_once.py msm2 20 0 > $O/${T}_g2_ncu.log 2>&1; echo "g2 launches $?"
python tools/launch_summary.py $O/${T}_g2_launches.csv > $O/${T}_g2_launches_summary.txt
REPS=1 timeout 900 $F -k regex:"msm_accumulate|ba_level|msm_halve|msm_reduce" -c 14 -o $O/${T}_g2_full python tools/run_once.py msm2 20 0 > $O/${T}_g2_full.log 2>&1; echo "g2 full $?"; sumrep ${T}_g2_full
REPS=1 timeout 300 $L -c 600 --log-file $O/${T}_msm16_launches.csv python tools/run_once.py msm 16 0 > $O/${T}_msm16_ncu.log 2>&1; echo "msm16 launches $?"
python tools/launch_summary.py $O/${T}_msm16_launches.csv > $O/${T}_msm16_launches_summary.txt
REPS=1 timeout 600 $F -k regex:"spmv|pointwise|finalize" -c 5 -o $O/${T}_prove_full python tools/run_once.py prove 20 0 > $O/${T}_prove_full.log 2>&1; echo "prove full $?"; sumrep ${T}_prove_full
REPS=1 timeout 600 $F -k regex:"fixed_base|lagrange|qap_columns|abc_kernel" -c 8 -o $O/${T}_setup_full python tools/run_once.py setup 18 > $O/${T}_setup_full.log 2>&1; echo "setup full $?"; sumrep ${T}_setup_full
REPS=1 timeout 600 $F -k regex:"te_mul" -c 2 -o $O/${T}_te_full python tools/run_once.py te 20 > $O/${T}_te_full.log 2>&1; echo "te full $?"; sumrep ${T}_te_full
du -sh $O; ls -la $O | tail -20
