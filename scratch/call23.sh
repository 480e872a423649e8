for t in 0 6 7; do GUDA_SIMT_TILE=$t timeout 300 python scratch/ab_tile.py 512 1024 1280 1536 2048; done > gpurun_out/r1u_ab.log 2>&1
cat gpurun_out/r1u_ab.log
(GUDA_SIMT_TILE=6 timeout 900 python -m pytest tests/test_gpu_gemm.py -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r1u_pytest6.log 2>&1
(GUDA_SIMT_TILE=7 timeout 900 python -m pytest tests/test_gpu_gemm.py -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r1u_pytest7.log 2>&1
cat gpurun_out/r1u_pytest6.log gpurun_out/r1u_pytest7.log
