(timeout 300 python -m pytest tests/test_gpu_tensor.py -m gpu -x -q 2>&1 | tail -15)
for v in 1 0 1 0; do echo "== GUDA_TC_PAIR=$v"; GUDA_TC_PAIR=$v timeout 120 python scratch/quick_bf16.py 2>&1 | tail -4; done
