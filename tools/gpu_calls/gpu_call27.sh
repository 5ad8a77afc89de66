O=gpurun_out/c27; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
python benchmarks/rep_bench.py 2> $O/rep.err > $O/rep_bench.jsonl; cut -c1-120,190-330 $O/rep_bench.jsonl
python benchmarks/secondary.py --cases c1,c1b 2>> $O/rep.err | tee $O/secondary_c1.jsonl | cut -c1-200
ncu --set full --clock-control none --import-source on -k regex:rep_trunk -s 3 -c 1 -o $O/rep_trunk_full python benchmarks/rep_bench.py > $O/ncu_rep.log 2>&1; echo "ncu rep rc=$?"
ncu -i $O/rep_trunk_full.ncu-rep --page raw --csv > $O/rep_trunk_full_raw.csv 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:knn_prep -s 3 -c 1 -o $O/prep_full python benchmarks/rep_bench.py > $O/ncu_prep.log 2>&1; echo "ncu prep rc=$?"
ncu -i $O/prep_full.ncu-rep --page raw --csv > $O/prep_full_raw.csv 2>/dev/null
rm -f $O/*.ncu-rep
tail -3 $O/rep.err
