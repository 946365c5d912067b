O=gpurun_out; T=r2d
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${T}_smi.txt 2>&1; nproc >> $O/${T}_smi.txt
(time timeout 1500 python -m pytest tests -m gpu -q --timeout 300 --durations=25) > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"
tail -45 $O/${T}_pytest.log
timeout 600 python bench.py --no-extras > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"; head -c 2500 $O/${T}_bench.json; tail -3 $O/${T}_bench.err
