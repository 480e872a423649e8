(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/r1f_pytest.log 2>&1
python scratch/quick_sgemm.py > gpurun_out/r1f_quick.log 2>&1
python bench.py --steps 1 --warmup 1 --no-suite --no-e2e --no-cpu > gpurun_out/r1f_plain_head.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_simt -s 8 -c 8 --csv --log-file gpurun_out/r1f_headline_traffic.csv python bench.py --steps 1 --warmup 1 --no-suite --no-e2e --no-cpu > gpurun_out/r1f_ncu_head.log 2>&1
cat gpurun_out/r1f_pytest.log gpurun_out/r1f_quick.log; cut -c1-200 gpurun_out/r1f_plain_head.log
