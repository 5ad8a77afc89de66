O=gpurun_out/c22; mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; tail -1 $O/bench_n1.json | cut -c1-600
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_ref.err; echo "ref rc=$?"; tail -1 $O/bench_reference.json | cut -c1-400
python benchmarks/secondary.py 2> $O/sec.err | tee $O/secondary.jsonl | cut -c1-260
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_bench_n1.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_bench.log 2>&1; echo "ncu rc=$?"
