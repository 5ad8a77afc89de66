O=gpurun_out/c38; mkdir -p $O
python -m pytest tests/test_gpu_relabel.py -x -q -k "c5_full_size" --durations=3 > $O/pytest_c5.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_c5.log
