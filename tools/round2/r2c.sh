O=gpurun_out; T=r2c
timeout 600 python -m pytest tests/test_gpu_multi_device.py -m gpu -x -q --timeout 120 > $O/${T}_pytest_multi.log 2>&1; echo "pytest-multi exit $?"
tail -30 $O/${T}_pytest_multi.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 200 > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -30 $O/${T}_pytest.log
