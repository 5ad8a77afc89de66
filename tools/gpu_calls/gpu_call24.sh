O=gpurun_out/c24; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu.log
python benchmarks/rep_bench.py 2> $O/rep.err > $O/rep_bench_r8.jsonl; cut -c1-120,190-330 $O/rep_bench_r8.jsonl
EXB_TRUNK_R=4 python benchmarks/rep_bench.py 2>> $O/rep.err > $O/rep_bench_r4.jsonl; cut -c1-120,190-330 $O/rep_bench_r4.jsonl
tail -3 $O/rep.err
