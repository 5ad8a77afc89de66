O=gpurun_out/c7; mkdir -p $O
python -m pytest tests/test_gpu_knn.py tests/test_gpu_relabel.py -m gpu -x -q > $O/pytest_knn.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_knn.log
python benchmarks/screen_scaling.py > $O/scaling_f3.jsonl 2> $O/err3.log; echo rc=$?; cat $O/scaling_f3.jsonl
EXB_SCREEN_TRACE=1 python benchmarks/screen_scaling.py --shards 1,8 --reps 1 > $O/scaling_f3_trace.jsonl 2> $O/err3t.log; echo rc=$?; cat $O/scaling_f3_trace.jsonl
EXB_SCREEN_FILTER=2 python benchmarks/screen_scaling.py --shards 1,8 > $O/scaling_f2.jsonl 2> $O/err2.log; echo rc=$?; cat $O/scaling_f2.jsonl
