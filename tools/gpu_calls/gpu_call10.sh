O=gpurun_out/c10; mkdir -p $O
python -m pytest tests/test_gpu_knn.py tests/test_gpu_relabel.py -m gpu -x -q > $O/pytest_knn.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_knn.log
for F in 3 2; do for H in 1 0; do
echo "FILTER=$F HEAD=$H"
EXB_SCREEN_FILTER=$F EXB_SCREEN_HEAD=$H python benchmarks/screen_scaling.py --shards 1,8 2> $O/e.log | tee $O/f${F}_h${H}.jsonl
done; done
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum,sm__inst_executed.sum,sm__cycles_elapsed.avg.per_second,l1tex__t_bytes_pipe_lsu_mem_local_op_ld.sum
for F in 3 2; do
EXB_SCREEN_FILTER=$F ncu --metrics $M --clock-control none -k regex:knn_screen_kernel -c 2 --csv --log-file $O/ncu_f$F.csv python benchmarks/screen_scaling.py --shards 2 --reps 1 > $O/run_f$F.log 2>&1; echo "rc=$?"
done
