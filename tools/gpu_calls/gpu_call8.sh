O=gpurun_out/c8; mkdir -p $O
python benchmarks/screen_scaling.py --shards 1,8 > $O/f3.jsonl 2> $O/e1.log; cat $O/f3.jsonl
EXB_SCREEN_SLEEP=200 python benchmarks/screen_scaling.py --shards 1,8 > $O/f3s.jsonl 2> $O/e2.log; cat $O/f3s.jsonl
EXB_SCREEN_FILTER=2 python benchmarks/screen_scaling.py --shards 1,8 > $O/f2.jsonl 2> $O/e3.log; cat $O/f2.jsonl
EXB_SCREEN_DEBUG=1 python benchmarks/screen_scaling.py --shards 1 > $O/d1.jsonl 2> $O/e4.log; cat $O/d1.jsonl
