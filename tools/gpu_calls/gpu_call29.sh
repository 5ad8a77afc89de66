O=gpurun_out/c29; mkdir -p $O
python -m pytest tests/test_gpu_rep.py -x -q > $O/pytest_rep.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_rep.log
python benchmarks/secondary.py --cases c1b,c3 2> $O/sec.err | tee $O/secondary_c1b_c3.jsonl | cut -c1-330 | grep -v "walker-state gather"
python benchmarks/rep_bench.py 2>> $O/sec.err > $O/rep_bench.jsonl; cut -c1-120,190-330 $O/rep_bench.jsonl | head -2
tail -3 $O/sec.err
