O=gpurun_out/c3; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$TR benchmarks/relabel_nccl_check.py --n 20000 > $O/nccl_check.log 2>&1; echo "nccl check rc=$?"; tail -5 $O/nccl_check.log
$TR bench.py --gpus 2 --steps 5 --warmup 3 --stage-times > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench2 rc=$?"; cat $O/bench_n2.json; grep "stage times" $O/bench_n2.err
$TR bench.py --gpus 2 --steps 5 --warmup 3 --exchange all_gather --stage-times > $O/bench_n2_ag.json 2> $O/bench_n2_ag.err; echo "bench2ag rc=$?"; cat $O/bench_n2_ag.json; grep "stage times" $O/bench_n2_ag.err
python benchmarks/secondary.py --cases c4 > $O/secondary_c4.jsonl 2>&1; cat $O/secondary_c4.jsonl
nvidia-smi topo -m > $O/topo.txt 2>&1
