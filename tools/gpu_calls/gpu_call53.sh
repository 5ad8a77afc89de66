O=gpurun_out/c53; mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"
python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_gpu.log
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; python tools/bench_line.py < $O/bench_n1.json
python benchmarks/secondary.py 2> $O/sec.err > $O/secondary.jsonl; echo "secondary rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_bench_n1.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_bench.log 2>&1; echo "ncu rc=$?"
