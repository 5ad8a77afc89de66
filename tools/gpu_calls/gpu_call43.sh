O=gpurun_out/c43; mkdir -p $O
for R in 1 0 1 0; do
EXB_SCREEN_CHAINRESET=$R python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $O/bench_r$R.json 2> $O/bench.err; echo -n "reset=$R "; python tools/bench_line.py < $O/bench_r$R.json
done
for R in 1 0; do
EXB_SCREEN_CHAINRESET=$R python benchmarks/screen_scaling.py --shards 8 --reps 4 2>> $O/e.log | cut -c1-60,100-260
EXB_SCREEN_CHAINRESET=$R python benchmarks/screen_scaling.py --shards 8 --reps 4 --query-blocks 2>> $O/e.log | cut -c1-60,100-260
done
