(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > gpurun_out/r1z_pytest.log 2>&1
cat gpurun_out/r1z_pytest.log
(timeout 300 python -c "import __graft_entry__ as e; e.smoke()" 2>&1 | tail -2) > gpurun_out/r1z_smoke.log 2>&1
cat gpurun_out/r1z_smoke.log
(time timeout 1200 python bench.py) > gpurun_out/r1z_bench.json 2> gpurun_out/r1z_bench.err
tail -c 600 gpurun_out/r1z_bench.json; tail -4 gpurun_out/r1z_bench.err
timeout 900 python bench.py --steps 2 --warmup 1 > gpurun_out/r1z_bench_s2.json 2> gpurun_out/r1z_bench_s2.err && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1z_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r1z_ncu.log 2>&1
tail -2 gpurun_out/r1z_ncu.log
