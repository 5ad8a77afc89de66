O=gpurun_out/c21; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
python benchmarks/secondary.py --cases c1,c2 2> $O/e.log | tee $O/secondary_c12.jsonl | cut -c1-260
python benchmarks/screen_scaling.py --shards 1,8 --reps 2 2>> $O/e.log | tee $O/scaling.jsonl | cut -c1-200
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_c1.csv python benchmarks/c1_once.py > $O/ncu_c1.log 2>&1; echo "ncu rc=$?"
grep -h "knn_prep\|knn_screen_lat\|rerank\|pbe_finish" $O/launches_c1.csv | cut -d, -f5,15 | head -12
