O=gpurun_out/c38; mkdir -p $O
python -m pytest tests/test_gpu_relabel.py tests/test_gpu_knn.py -x -q -k "hybrid or randomized or outside_the_screen or query_block" > $O/pytest_new.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_new.log
