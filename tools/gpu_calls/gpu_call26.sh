G=${1:-2}
O=gpurun_out/c26_$G; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511"
$TR benchmarks/relabel_nccl_check.py > $O/nccl_check.log 2>&1; echo "nccl check rc=$?"; grep "nccl check" $O/nccl_check.log; tail -3 $O/nccl_check.log | cut -c1-300
$TR bench.py --gpus $G --steps 5 --warmup 3 --stage-times > $O/bench_auto.json 2> $O/bench_auto.err; echo "bench auto rc=$?"; tail -1 $O/bench_auto.json | cut -c1-1200; grep "stage times" $O/bench_auto.err
$TR bench.py --gpus $G --steps 5 --warmup 3 --exchange lists --stage-times > $O/bench_lists.json 2> $O/bench_lists.err; echo "bench lists rc=$?"; tail -1 $O/bench_lists.json | cut -c1-330; grep "stage times" $O/bench_lists.err
