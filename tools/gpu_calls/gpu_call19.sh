O=gpurun_out/c19; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$TR benchmarks/relabel_nccl_check.py > $O/nccl_check.log 2>&1; echo "nccl check rc=$?"; grep "nccl check\|Error\|error" $O/nccl_check.log | head
for X in auto all_to_all; do
$TR bench.py --gpus 2 --steps 5 --warmup 3 --exchange $X --stage-times > $O/bench_n2_$X.json 2> $O/bench_n2_$X.err; echo "bench2 $X rc=$?"; python -c "
import json,sys; d=json.load(open('$O/bench_n2_$X.json')); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['config']['parallelism'])"; grep "stage times" $O/bench_n2_$X.err
done
