(timeout 900 python -m pytest tests/test_gpu_host_gemm.py -m gpu -x -q 2>&1 | tail -8) > gpurun_out/r1j_pytest.log 2>&1
cat gpurun_out/r1j_pytest.log
