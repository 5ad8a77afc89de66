O=gpurun_out/c44; mkdir -p $O
python -m pytest tests/test_gpu_relabel.py tests/test_gpu_knn.py tests/test_gpu_rep.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
python benchmarks/trajectory_data_check.py 2> $O/e.log | tee $O/trajectory.jsonl | cut -c1-330
tail -2 $O/e.log
