export GUDA_GEMVT_COLS=1
python scratch/prof_gemvt.py && ncu --set full --clock-control none --import-source on -k regex:sgemv_t -s 2 -c 1 -o gpurun_out/r1g_gemvt_ws python scratch/prof_gemvt.py > gpurun_out/r1g_ncu.log 2>&1
tail -3 gpurun_out/r1g_ncu.log
