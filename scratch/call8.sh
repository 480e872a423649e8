(timeout 900 python -m pytest tests/test_gpu_level2.py -m gpu -x -q 2>&1 | tail -12) > gpurun_out/r1g_pytest.log 2>&1
timeout 300 python scratch/quick_l2.py > gpurun_out/r1g_l2.log 2>&1
cat gpurun_out/r1g_pytest.log gpurun_out/r1g_l2.log
