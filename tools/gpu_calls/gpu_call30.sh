O=gpurun_out/c30; mkdir -p $O
ncu --metrics gpu__time_duration.sum,smsp__cycles_active.avg,sm__cycles_active.avg,launch__grid_size,smsp__inst_executed.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,dram__bytes_read.sum,launch__occupancy_limit_shared_mem,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"rep_trunk" -c 6 --csv --log-file $O/launches_trunk.csv python benchmarks/secondary.py --cases c1b > $O/ncu_c1b.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open("gpurun_out/c30/launches_trunk.csv")) if len(r)>10]
hdr=rows[0]
ki=hdr.index("Kernel Name"); mi=hdr.index("Metric Name"); vi=hdr.index("Metric Value"); ii=hdr.index("ID")
seen={}
for r in rows[1:]:
    seen.setdefault(r[ii],{"k":r[ki][:40]})[r[mi]]=r[vi]
for i,(k,v) in enumerate(seen.items()):
    if i in (0,3,5): print(v)
PY
