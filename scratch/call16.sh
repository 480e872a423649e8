(timeout 900 python -m pytest tests/test_gpu_conv.py -m gpu -x -q 2>&1 | tail -12) > gpurun_out/r1n_pytest_conv.log 2>&1
cat gpurun_out/r1n_pytest_conv.log
timeout 300 python scratch/quick_conv.py > gpurun_out/r1n_conv.log 2>&1
cat gpurun_out/r1n_conv.log
