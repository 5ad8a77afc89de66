O=gpurun_out/c14; mkdir -p $O
for S in 1 2 8 32; do echo "SEED=$S"; EXB_SCREEN_SEED=$S python benchmarks/screen_scaling.py --shards 1,8 2>> $O/e.log | tee $O/f3_seed$S.jsonl; done
echo F2; EXB_SCREEN_FILTER=2 python benchmarks/screen_scaling.py --shards 1,8 2>> $O/e.log | tee $O/f2.jsonl
for S in 1 32; do echo "SEED=$S again"; EXB_SCREEN_SEED=$S python benchmarks/screen_scaling.py --shards 1,8 2>> $O/e.log | tee $O/f3_seed${S}b.jsonl; done
