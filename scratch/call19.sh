timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_simt -c 2 -o gpurun_out/r1q_sgemm4096 -f python scratch/ab_tile.py 4096 > gpurun_out/r1q_ncu.log 2>&1
tail -3 gpurun_out/r1q_ncu.log
