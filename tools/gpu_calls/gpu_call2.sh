O=gpurun_out/c2; mkdir -p $O
python -m pytest tests/test_gpu_replay.py -m gpu -x -q > $O/pytest_replay.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_replay.log
python benchmarks/secondary.py --cases c3,c4 > $O/secondary_c34.jsonl 2> $O/secondary.err; echo "secondary rc=$?"; cat $O/secondary_c34.jsonl
ncu --set full --clock-control none --import-source on -k regex:replay_gather -s 12 -c 1 -o $O/gather_full -f python benchmarks/secondary.py --cases c4 > $O/ncu_gather.log 2>&1; echo "ncu rc=$?"
