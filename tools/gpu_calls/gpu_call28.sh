O=gpurun_out/c28; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
python benchmarks/secondary.py --cases c1b,c3 2> $O/sec.err | tee $O/secondary_c1b_c3.jsonl | cut -c1-420
tail -3 $O/sec.err
