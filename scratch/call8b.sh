timeout 300 python scratch/quick_l2.py > gpurun_out/r1g_l2.log 2>&1
grep "sgemv T" gpurun_out/r1g_l2.log
