set -x
python bench.py --steps 2 --warmup 1 > gpurun_out/r1_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_bench.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r1_ncu_bench.log 2>&1
PROF_REPS=1 python profiles/prof_kernels.py all > gpurun_out/r1_plain_prof.log 2>&1 && \
PROF_REPS=1 ncu --set full --clock-control none --import-source on -k regex:'axpy|dot_kernel|scal|gemm_simt|gemm_tc|dgemm' -o gpurun_out/r1_prof_all python profiles/prof_kernels.py all > gpurun_out/r1_ncu_prof.log 2>&1
tail -3 gpurun_out/r1_ncu_bench.log gpurun_out/r1_ncu_prof.log
ls -la gpurun_out
