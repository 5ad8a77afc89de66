O=gpurun_out/c49; mkdir -p $O
python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
python benchmarks/secondary.py --cases c1,c2 2> $O/sec.err > $O/secondary.jsonl; echo "secondary rc=$?"; cut -c1-200 $O/secondary.jsonl
python benchmarks/trajectory_data_check.py --per-update-only > $O/trajectory.jsonl 2> $O/traj.err; echo "traj rc=$?"; cat $O/trajectory.jsonl
