O=gpurun_out; T=r2z3
mkdir -p $O
(time timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300) > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
for w in "poe 1" "poe 100" "prove 8" "prove 12" "prove 14"; do echo "$w"; REPS=9 timeout 300 python tools/run_once.py $w 0 2>&1 | grep total_ms | sed "s/.*'total_ms': \([0-9.]*\).*/\1/" | sort -n | head -5 | tr '\n' ' '; echo; done
(time python -c "import __graft_entry__ as g; g.smoke()") 2>&1 | tail -4
