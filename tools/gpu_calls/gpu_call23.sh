O=gpurun_out/c23; mkdir -p $O
python -m pytest tests/test_gpu_rep.py -x -q > $O/pytest_rep.log 2>&1; echo "pytest rep rc=$?"; tail -15 $O/pytest_rep.log
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
python benchmarks/rep_bench.py 2> $O/rep.err | tee $O/rep_bench_r8.jsonl | cut -c1-330
EXB_TRUNK_R=4 python benchmarks/rep_bench.py 2>> $O/rep.err | tee $O/rep_bench_r4.jsonl | head -2 | cut -c1-330
tail -3 $O/rep.err
ncu --set full --clock-control none --import-source on -k regex:rep_trunk -s 3 -c 1 -o $O/rep_trunk_full python benchmarks/rep_bench.py > $O/ncu_rep.log 2>&1; echo "ncu rep rc=$?"
ncu -i $O/rep_trunk_full.ncu-rep --page raw --csv > $O/rep_trunk_full_raw.csv 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:knn_screen_kernel -s 3 -c 1 -o $O/screen_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_screen.log 2>&1; echo "ncu screen rc=$?"
ncu -i $O/screen_full.ncu-rep --page raw --csv > $O/screen_full_raw.csv 2>/dev/null
ls -la $O
