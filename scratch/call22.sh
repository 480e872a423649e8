(time timeout 1200 python -m pytest tests/test_gpu_full_size.py -m gpu -x -q --durations=6 2>&1 | tail -25) > gpurun_out/r1t_pytest_full.log 2>&1
cat gpurun_out/r1t_pytest_full.log
