N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 1 > gpurun_out/mgpu_bench_$N.log 2>&1
echo "bench rc=$?"; grep -c '"metric"' gpurun_out/mgpu_bench_$N.log
