for t in 0 1 2 3; do GUDA_SIMT_TILE=$t timeout 300 python scratch/ab_tile.py; done > gpurun_out/r1m_ab_tile.log 2>&1
cat gpurun_out/r1m_ab_tile.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_simt -c 6 -o gpurun_out/r1m_sgemm1024 -f python scratch/ab_tile.py 1024 > gpurun_out/r1m_ncu.log 2>&1
tail -3 gpurun_out/r1m_ncu.log
