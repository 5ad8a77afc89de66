O=gpurun_out/c40; mkdir -p $O
EXB_SCREEN_TRACE=1 python benchmarks/trajectory_data_check.py 2> $O/e.log | tee $O/trajectory.jsonl
tail -3 $O/e.log
