O=gpurun_out/c33; mkdir -p $O
python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
python benchmarks/secondary.py --cases c1,c1b,c2,c3 2> $O/sec.err | tee $O/secondary_pdl.jsonl | cut -c1-330 | grep -v "exact\|torch trunk\|walker-state gather"
EXB_PDL=0 python benchmarks/secondary.py --cases c1,c2 2>> $O/sec.err | tee $O/secondary_nopdl.jsonl | cut -c1-330 | grep -v "exact"
tail -3 $O/sec.err
