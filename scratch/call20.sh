for t in 0 4 2; do GUDA_SIMT_TILE=$t timeout 300 python scratch/ab_tile.py 512 768 1024 1536 2048 3072 4096; done > gpurun_out/r1r_ab_tile.log 2>&1
cat gpurun_out/r1r_ab_tile.log
(GUDA_SIMT_TILE=4 timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_host_gemm.py tests/test_gpu_conv.py -m gpu -x -q 2>&1 | tail -5) > gpurun_out/r1r_pytest_tile4.log 2>&1
cat gpurun_out/r1r_pytest_tile4.log
