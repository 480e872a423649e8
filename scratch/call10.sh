set -x
(timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -6) > gpurun_out/r1h_pytest.log 2>&1
python bench.py > gpurun_out/r1h_bench.json 2> gpurun_out/r1h_bench.err
python bench.py --steps 2 --warmup 1 > gpurun_out/r1h_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1h_launches_bench.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r1h_ncu_bench.log 2>&1
PROF_REPS=1 PROF_N=4096 python profiles/prof_kernels.py all > gpurun_out/r1h_plain_prof.log 2>&1 && \
PROF_REPS=1 PROF_N=4096 ncu --set full --clock-control none --import-source on -k regex:'axpy|dot_kernel|sgemv|sger|gemm_simt|gemm_tc|dgemm_dmma' -o gpurun_out/r1h_prof python profiles/prof_kernels.py all > gpurun_out/r1h_ncu_prof.log 2>&1
cat gpurun_out/r1h_pytest.log; cut -c1-600 gpurun_out/r1h_bench.json; tail -2 gpurun_out/r1h_ncu_prof.log
