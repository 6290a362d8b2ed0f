# ncu source-level counters of the framed integer AVG kernel (one launch)
mkdir -p gpurun_out
REP=/tmp/r2_window3
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'an_frame_kernel' \
  --launch-skip 2 --launch-count 1 -o $REP -f python bench.py --workload window --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_window3.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_window3.log
ncu -i $REP.ncu-rep --page source --csv > gpurun_out/r2_window3_avg_source.csv 2>/dev/null
ls -la gpurun_out/r2_window3_avg_source.csv
