O=gpurun_out/c16; mkdir -p $O
python -m pytest tests/test_gpu_knn.py tests/test_gpu_relabel.py -m gpu -x -q > $O/pytest_knn.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_knn.log
for F in 2 3; do for S in -1 0 32; do echo "FILTER=$F SEED=$S"; EXB_SCREEN_FILTER=$F EXB_SCREEN_SEED=$S python benchmarks/screen_scaling.py --shards 1,2,8 2>> $O/e.log | tee $O/f${F}_seed$S.jsonl; done; done
