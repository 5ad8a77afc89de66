O=gpurun_out/c42; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_gpu.log
EXB_SCREEN_TRACE=1 python benchmarks/trajectory_data_check.py 2> $O/e.log | tee $O/trajectory.jsonl | cut -c1-420
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; python tools/bench_line.py < $O/bench_n1.json
python benchmarks/screen_scaling.py --shards 8 --reps 4 2>> $O/e.log | cut -c1-60,100-300
python benchmarks/secondary.py --cases c1,c2 2>> $O/e.log | grep screened | cut -c1-200
