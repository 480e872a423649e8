for t in 0 5 0 5; do GUDA_SIMT_TILE=$t timeout 300 python scratch/ab_tile.py 1536 2048 4096 8192; done > gpurun_out/r1s_ab_tall.log 2>&1
cat gpurun_out/r1s_ab_tall.log
(GUDA_SIMT_TILE=5 timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_host_gemm.py -m gpu -x -q 2>&1 | tail -5) > gpurun_out/r1s_pytest_tile5.log 2>&1
cat gpurun_out/r1s_pytest_tile5.log
