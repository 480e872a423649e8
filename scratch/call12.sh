(timeout 900 python -m pytest tests/test_gpu_host_gemm.py -m gpu -x -q 2>&1 | tail -4) > gpurun_out/r1j_pytest.log 2>&1
cat gpurun_out/r1j_pytest.log
python bench.py --steps 1 --warmup 1 --no-suite --no-cpu > gpurun_out/r1j_bench.json 2> gpurun_out/r1j_bench.err; tail -3 gpurun_out/r1j_bench.err
python -c "
import json; d=json.loads(open('gpurun_out/r1j_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e'])"
