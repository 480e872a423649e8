timeout 300 python scratch/prof_late.py && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ew_kernel|fused_execute|reduce_axis|sgemv_t_kernel|conv_igemm|gemm_simt' --launch-skip 0 -c 40 -o gpurun_out/r2c_late -f python scratch/prof_late.py > gpurun_out/r2c_ncu.log 2>&1
tail -2 gpurun_out/r2c_ncu.log
