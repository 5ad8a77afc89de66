mkdir -p gpurun_out/c1
O=gpurun_out/c1
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"
python bench.py --stage-times > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
python benchmarks/secondary.py > $O/secondary.jsonl 2> $O/secondary.err; echo "secondary rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_bench_n1.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_bench.log 2>&1; echo "ncu1 rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c1.csv python benchmarks/c1_once.py > $O/ncu_c1.log 2>&1; echo "ncu2 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:replay_gather -s 12 -c 1 -o $O/gather_full -f python benchmarks/secondary.py --cases c4 --c4-transitions 200000 > $O/ncu_gather.log 2>&1; echo "ncu3 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'pbe_|knn_rerank|knn_prep' -s 12 -c 4 -o $O/epi_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_epi.log 2>&1; echo "ncu4 rc=$?"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt
