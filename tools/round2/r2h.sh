O=gpurun_out; T=r2h
mkdir -p $O
F="ncu --set full --clock-control none --import-source on -f"
REPS=1 timeout 600 $F -k regex:"ntt_pass" -s 0 -c 3 -o $O/${T}_ntt python tools/run_once.py ntt 22 > $O/${T}_ntt.log 2>&1; echo "ncu $?"
python tools/ncu_summary.py $O/${T}_ntt.ncu-rep > $O/${T}_ntt_summary.txt 2>&1
ncu -i $O/${T}_ntt.ncu-rep --page details --csv 2>/dev/null | grep -i "stall\|Warp Cycles Per\|Issued Warp\|Eligible\|No Eligible\|Active Warps\|One or More" | head -60 > $O/${T}_ntt_details.txt
ncu -i $O/${T}_ntt.ncu-rep --page raw --csv --metrics regex:"smsp__average_warp.*_per_issue_active|smsp__pcsamp_warps_issue_stalled.*" 2>/dev/null > $O/${T}_ntt_stalls.csv
rm -f $O/${T}_ntt.ncu-rep
ls -la $O | grep r2h
