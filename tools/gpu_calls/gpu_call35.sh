O=gpurun_out/c35; mkdir -p $O
python -m pytest tests/test_gpu_relabel.py tests/test_gpu_knn.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
for T in 1 0; do
EXB_SCREEN_TAIL=$T python benchmarks/screen_scaling.py --shards 8,4 --reps 6 --query-blocks 2>> $O/e.log | tee -a $O/qblocks_tail$T.jsonl | cut -c1-60,100-330
done
tail -3 $O/e.log
