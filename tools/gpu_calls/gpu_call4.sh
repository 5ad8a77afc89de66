O=gpurun_out/c4; mkdir -p $O
python benchmarks/screen_scaling.py > $O/scaling_d0.jsonl 2> $O/err0.log; echo rc=$?; cat $O/scaling_d0.jsonl
EXB_SCREEN_DEBUG=1 python benchmarks/screen_scaling.py > $O/scaling_d1.jsonl 2> $O/err1.log; echo rc=$?; cat $O/scaling_d1.jsonl
