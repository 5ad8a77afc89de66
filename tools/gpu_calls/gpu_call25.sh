O=gpurun_out/c25; mkdir -p $O
python -m pytest tests/test_gpu_rep.py tests/test_gpu_relabel.py -x -q > $O/pytest_new.log 2>&1; echo "pytest new rc=$?"; tail -5 $O/pytest_new.log
python benchmarks/screen_scaling.py --shards 8 --reps 3 --query-blocks 2> $O/e.log | tee $O/qblocks_auto.jsonl | cut -c1-300
EXB_SCREEN_WAVESPLITS=1 python benchmarks/screen_scaling.py --shards 8 --reps 3 --query-blocks 2>> $O/e.log | tee $O/qblocks_s1.jsonl | cut -c1-300
EXB_SCREEN_WAVESPLITS=2 python benchmarks/screen_scaling.py --shards 8 --reps 3 --query-blocks 2>> $O/e.log | tee $O/qblocks_s2.jsonl | cut -c1-300
EXB_SCREEN_WAVESPLITS=4 python benchmarks/screen_scaling.py --shards 8 --reps 3 --query-blocks 2>> $O/e.log | tee $O/qblocks_s4.jsonl | cut -c1-300
python benchmarks/screen_scaling.py --shards 2,4 --reps 2 --query-blocks 2>> $O/e.log | tee $O/qblocks_24.jsonl | cut -c1-300
python benchmarks/screen_scaling.py --shards 8 --reps 3 2>> $O/e.log | tee $O/bankshard8.jsonl | cut -c1-300
python benchmarks/rep_bench.py 2>> $O/e.log > $O/rep_bench.jsonl; cut -c1-120,190-330 $O/rep_bench.jsonl | head -2
tail -3 $O/e.log
