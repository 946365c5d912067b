O=gpurun_out; T=r2z5
mkdir -p $O
(time timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300) > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
(time python -c "import __graft_entry__ as g; g.smoke()") 2>&1 | tail -4
timeout 600 python bench.py --no-extras > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"; head -c 700 $O/${T}_bench.json
