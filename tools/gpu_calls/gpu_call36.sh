G=${1:-8}
O=gpurun_out/c36_$G; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $G --steps 5 --warmup 3 --stage-times > $O/bench_auto.json 2> $O/bench_auto.err; echo "bench auto rc=$?"; tail -1 $O/bench_auto.json | cut -c1-330; grep "stage times" $O/bench_auto.err
