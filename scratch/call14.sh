(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/r1l_pytest.log 2>&1
cat gpurun_out/r1l_pytest.log
(timeout 300 python -c "import __graft_entry__ as e; e.smoke()" 2>&1 | tail -3) > gpurun_out/r1l_smoke.log 2>&1
cat gpurun_out/r1l_smoke.log
(time timeout 1200 python bench.py) > gpurun_out/r1l_bench.json 2> gpurun_out/r1l_bench.err
tail -c 3000 gpurun_out/r1l_bench.json; tail -5 gpurun_out/r1l_bench.err
(time timeout 600 python bench.py --impl reference --steps 2 --warmup 1) > gpurun_out/r1l_bench_ref.json 2> gpurun_out/r1l_bench_ref.err
cat gpurun_out/r1l_bench_ref.json; tail -4 gpurun_out/r1l_bench_ref.err
