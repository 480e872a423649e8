(time timeout 1200 python bench.py) > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
tail -c 300 gpurun_out/r2b_bench.json; tail -4 gpurun_out/r2b_bench.err
(timeout 300 python -c "import __graft_entry__ as e; e.smoke()" 2>&1 | tail -2)
