O=gpurun_out/c13; mkdir -p $O
python -m pytest tests/test_gpu_knn.py tests/test_gpu_relabel.py -m gpu -x -q > $O/pytest_knn.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_knn.log
python benchmarks/screen_scaling.py --shards 1,2,4,8 2> $O/e.log | tee $O/f3.jsonl
for S in 4 16 32; do echo "SEED=$S"; EXB_SCREEN_SEED=$S python benchmarks/screen_scaling.py --shards 1,8 2>> $O/e.log | tee $O/f3_seed$S.jsonl; done
EXORL_B200_LIB=$PWD/exorl_b200/lib/libexorl_b200_stats.so EXB_SCREEN_TRACE=1 python benchmarks/screen_scaling.py --shards 1,8 --reps 1 2>> $O/e.log | tee $O/f3_trace.jsonl
