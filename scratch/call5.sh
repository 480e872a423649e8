set -x
PROF_REPS=1 PROF_N=4096 python profiles/prof_kernels.py all > gpurun_out/r1d_plain_prof.log 2>&1 && \
PROF_REPS=1 PROF_N=4096 ncu --set full --clock-control none --import-source on -k regex:'axpy|dot_kernel|gemm_simt|gemm_tc|dgemm_dmma' -o gpurun_out/r1d_prof python profiles/prof_kernels.py all > gpurun_out/r1d_ncu_prof.log 2>&1
python bench.py --steps 1 --warmup 1 --no-suite --no-e2e --no-cpu > gpurun_out/r1d_plain_head.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_simt -c 2 --csv --log-file gpurun_out/r1d_headline_traffic.csv python bench.py --steps 1 --warmup 1 --no-suite --no-e2e --no-cpu > gpurun_out/r1d_ncu_head.log 2>&1
cat gpurun_out/r1d_plain_head.log | cut -c1-400
