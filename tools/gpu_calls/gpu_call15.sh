O=gpurun_out/c15; mkdir -p $O
python -m pytest tests/test_gpu_knn.py tests/test_gpu_relabel.py -m gpu -x -q > $O/pytest_knn.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_knn.log
echo F2; EXB_SCREEN_FILTER=2 python benchmarks/screen_scaling.py --shards 1,8 2>> $O/e.log | tee $O/f2.jsonl
for S in 2 32; do for C in 64 24; do echo "SEED=$S CAP=$C"; EXB_SCREEN_SEED=$S EXB_SCREEN_SCANCAP=$C python benchmarks/screen_scaling.py --shards 1,8 2>> $O/e.log | tee $O/f3_seed${S}_cap$C.jsonl; done; done
echo F2; EXB_SCREEN_FILTER=2 python benchmarks/screen_scaling.py --shards 1,8 2>> $O/e.log | tee $O/f2b.jsonl
