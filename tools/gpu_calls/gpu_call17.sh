O=gpurun_out/c17; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_gpu.log
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; cat $O/bench_n1.json
python benchmarks/secondary.py > $O/secondary.jsonl 2> $O/secondary.err; echo "secondary rc=$?"; cat $O/secondary.jsonl
python benchmarks/screen_scaling.py --shards 1,2,4,8 2> $O/e.log | tee $O/scaling.jsonl
