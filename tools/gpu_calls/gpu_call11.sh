O=gpurun_out/c11; mkdir -p $O
EXORL_B200_LIB=$PWD/exorl_b200/lib/libexorl_b200_stats.so EXB_SCREEN_TRACE=1 python benchmarks/screen_scaling.py --shards 1,8 --reps 1 2> $O/e.log | tee $O/f3_trace.jsonl
for S in 0 500 2000; do echo "SLEEP=$S"; EXB_SCREEN_SLEEP=$S python benchmarks/screen_scaling.py --shards 1,8 2>> $O/e.log | tee $O/f3_sleep$S.jsonl; done
