O=gpurun_out/c5; mkdir -p $O
python -m pytest tests/test_gpu_knn.py tests/test_gpu_relabel.py -m gpu -x -q > $O/pytest_knn.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_knn.log
python benchmarks/screen_scaling.py > $O/scaling_f2.jsonl 2> $O/err2.log; echo rc=$?; cat $O/scaling_f2.jsonl
EXB_SCREEN_FILTER=0 python benchmarks/screen_scaling.py --shards 1,8 > $O/scaling_f0.jsonl 2> $O/err0.log; echo rc=$?; cat $O/scaling_f0.jsonl
