#!/bin/bash
# One GPU-box session: parity tests, bench (ours + reference arm), ncu launch list and full captures.
# usage (from the repo root): gpurun --timeout 1800 -- 'bash tools/gpu_round.sh TAG'
TAG=${1:-r01b}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
   python bench.py --steps 2 --warmup 3 --no-extras > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:msm_accumulate -s 3 -c 1 -f -o $O/${TAG}_acc \
   python bench.py --steps 2 --warmup 3 --no-extras > $O/${TAG}_ncu_acc.log 2>&1; echo "ncu acc exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ntt -s 4 -c 4 -f -o $O/${TAG}_ntt \
   python tools/run_once.py ntt 22 > $O/${TAG}_ncu_ntt.log 2>&1; echo "ncu ntt exit $?"
ls -la $O | tail -20
