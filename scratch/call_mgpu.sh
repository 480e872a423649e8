N=${1:-2}
nvidia-smi -L
nvidia-smi topo -m 2>/dev/null | head -12
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py --size 16384 --iters 3 > gpurun_out/mgpu_check_$N.log 2>&1
echo "check rc=$?"; tail -5 gpurun_out/mgpu_check_$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 1 > gpurun_out/mgpu_bench_$N.log 2>&1
echo "bench rc=$?"; tail -3 gpurun_out/mgpu_bench_$N.log | cut -c1-1500
