O=gpurun_out; T=r2b
(time timeout 1200 python -m pytest tests/test_gpu_multi_device.py -m gpu -x -q --durations=8) > $O/${T}_pytest_multi.log 2>&1; echo "pytest-multi exit $?"
tail -30 $O/${T}_pytest_multi.log
(time timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q) > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $O/${T}_pytest.log
