(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > gpurun_out/r1v_pytest.log 2>&1
cat gpurun_out/r1v_pytest.log
for t in 0 4; do GUDA_SIMT_TILE=$t timeout 300 python scratch/ab_tile.py 512 768 1024 1280 1536 1792 2048 2560; done > gpurun_out/r1v_ab.log 2>&1
cat gpurun_out/r1v_ab.log
timeout 300 python scratch/quick_conv.py > gpurun_out/r1v_conv.log 2>&1
cat gpurun_out/r1v_conv.log
