O=gpurun_out/c47; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 5 --warmup 3 > $O/bench_n8.json 2> $O/bench_n8.err; echo "bench8 rc=$?"; python tools/bench_line.py < $O/bench_n8.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 benchmarks/relabel_nccl_check.py > $O/nccl_check_n8.log 2>&1; echo "nccl check rc=$?"; tail -3 $O/nccl_check_n8.log
