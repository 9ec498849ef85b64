(time timeout 900 python -m pytest tests -m gpu -x -q) > gpurun_out/r12_pytest.log 2>&1; tail -4 gpurun_out/r12_pytest.log
