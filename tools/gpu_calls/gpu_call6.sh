O=gpurun_out/c6; mkdir -p $O
python benchmarks/screen_scaling.py --shards 1,8 > $O/scaling_f2.jsonl 2> $O/err2.log; echo rc=$?; cat $O/scaling_f2.jsonl
EXB_SCREEN_TRACE=1 python benchmarks/screen_scaling.py --shards 1,8 --reps 1 > $O/scaling_trace.jsonl 2> $O/err3.log; echo rc=$?; cat $O/scaling_trace.jsonl
