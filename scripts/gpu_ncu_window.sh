# ncu --set full of the window kernels that changed this round (avg with prefix leaves, monotone RANGE search, prefix write)
mkdir -p gpurun_out
REP=/tmp/r2_window_full
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'an_frame_kernel|an_range_sorted_kernel|an_prefix_write_kernel|an_frame_prefix_kernel' \
  --launch-skip 8 --launch-count 7 -o $REP -f python bench.py --workload window --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_window.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_window.log
ncu -i $REP.ncu-rep --page details --csv > gpurun_out/r2_window_full_details.csv 2>/dev/null
ls -la gpurun_out/r2_window_full_details.csv
