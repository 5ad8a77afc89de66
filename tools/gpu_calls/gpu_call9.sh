O=gpurun_out/c9; mkdir -p $O
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum,sm__inst_executed.sum,sm__inst_executed_pipe_lsu.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,sm__cycles_elapsed.avg,sm__cycles_active.avg,smsp__thread_inst_executed.sum,l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum,l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum,l1tex__t_bytes_pipe_lsu_mem_local_op_ld.sum,l1tex__t_bytes_pipe_lsu_mem_local_op_st.sum,sm__cycles_elapsed.avg.per_second
for F in 3 2; do
EXB_SCREEN_FILTER=$F ncu --metrics $M --clock-control none -k regex:knn_screen_kernel -c 2 --csv --log-file $O/ncu_f$F.csv python benchmarks/screen_scaling.py --shards 2 --reps 1 > $O/run_f$F.log 2>&1; echo "rc=$?"
done
