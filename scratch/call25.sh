for t in 3 8 2; do GUDA_SIMT_TILE=$t timeout 300 python scratch/ab_tile.py 256 512 768 1024; done > gpurun_out/r1w_ab.log 2>&1
cat gpurun_out/r1w_ab.log
(GUDA_SIMT_TILE=8 timeout 900 python -m pytest tests/test_gpu_gemm.py -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r1w_pytest8.log 2>&1
cat gpurun_out/r1w_pytest8.log
