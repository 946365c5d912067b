O=gpurun_out; T=r2w
mkdir -p $O
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --timeout 200 > $O/${T}_pytest_alone.log 2>&1; echo "alone exit $?"; tail -4 $O/${T}_pytest_alone.log
(time timeout 1500 python -m pytest tests -m gpu -x -q --timeout 300) > $O/${T}_pytest_all.log 2>&1; echo "all exit $?"; tail -5 $O/${T}_pytest_all.log
