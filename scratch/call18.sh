timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -c 2 -o gpurun_out/r1p_conv -f python scratch/prof_conv.py > gpurun_out/r1p_ncu.log 2>&1
tail -3 gpurun_out/r1p_ncu.log
