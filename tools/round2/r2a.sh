O=gpurun_out; T=r2a
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${T}_smi.txt 2>&1; nproc >> $O/${T}_smi.txt
(time timeout 1200 python -m pytest tests/test_gpu_baseline_sizes.py tests/test_reference_constants.py -m gpu -x -q --durations=15) > $O/${T}_pytest_new.log 2>&1; echo "pytest-new exit $?"
tail -25 $O/${T}_pytest_new.log
(time timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_baseline_sizes.py) > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $O/${T}_pytest.log
timeout 600 python bench.py --no-extras > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"; cat $O/${T}_bench.json | head -c 3000
(time timeout 900 python bench.py --impl reference --steps 2 --warmup 1) > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref exit $?"; cat $O/${T}_bench_ref.json | head -c 1500; tail -4 $O/${T}_bench_ref.err
