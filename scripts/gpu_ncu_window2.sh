# ncu --set full of the staged RANGE search, the framed AVG and SUM kernels (one launch each, second step)
mkdir -p gpurun_out
REP=/tmp/r2_window2_full
timeout 500 ncu --set full --clock-control none --import-source on -k regex:'an_range_staged_kernel|an_frame_kernel|an_frame_prefix_kernel' \
  --launch-skip 6 --launch-count 4 -o $REP -f python bench.py --workload window --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_window2.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_window2.log
ncu -i $REP.ncu-rep --page details --csv > gpurun_out/r2_window2_full_details.csv 2>/dev/null
ncu -i $REP.ncu-rep --page source --csv --kernel-name regex:an_range_staged > gpurun_out/r2_window2_range_source.csv 2>/dev/null
ls -la gpurun_out/r2_window2_full_details.csv gpurun_out/r2_window2_range_source.csv
