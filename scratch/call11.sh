(timeout 900 python -m pytest tests/test_gpu_reduce.py -m gpu -x -q 2>&1 | tail -12) > gpurun_out/r1i_pytest.log 2>&1
cat gpurun_out/r1i_pytest.log
