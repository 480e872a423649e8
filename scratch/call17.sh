(timeout 900 python -m pytest tests/test_gpu_conv.py -m gpu -x -q 2>&1 | tail -5) > gpurun_out/r1o_pytest_conv.log 2>&1
cat gpurun_out/r1o_pytest_conv.log
timeout 300 python scratch/quick_conv.py > gpurun_out/r1o_conv.log 2>&1
cat gpurun_out/r1o_conv.log
for v in w162 w84 w162 w84; do echo "== $v"; GUDA_B200_LIB=$PWD/scratch/variants/lib$v.so timeout 300 python scratch/ab_tile.py 1024 2048 4096 8192; done > gpurun_out/r1o_ab_warp.log 2>&1
cat gpurun_out/r1o_ab_warp.log
(GUDA_B200_LIB=$PWD/scratch/variants/libw84.so timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_host_gemm.py -m gpu -x -q 2>&1 | tail -5) > gpurun_out/r1o_pytest_gemm_w84.log 2>&1
cat gpurun_out/r1o_pytest_gemm_w84.log
