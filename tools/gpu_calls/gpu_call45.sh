O=gpurun_out/c45; mkdir -p $O
python benchmarks/trajectory_data_check.py 2> $O/e.log | tee $O/trajectory.jsonl | cut -c1-330
tail -2 $O/e.log
