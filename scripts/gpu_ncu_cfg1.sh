mkdir -p gpurun_out
REP=/tmp/r2_cfg1_full
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'grp_hash_first|grp_assign_kernel|join_count|join_build' \
  --launch-skip 6 --launch-count 4 -o $REP -f python bench.py --workload config1 --steps 2 --warmup 3 > gpurun_out/ncu_cfg1.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_cfg1.log
ncu -i $REP.ncu-rep --page details --csv > gpurun_out/r2_cfg1_full_details.csv 2>/dev/null
ncu -i $REP.ncu-rep --page source --csv --print-source sass > gpurun_out/r2_cfg1_source.csv 2>/dev/null
ls -la gpurun_out/r2_cfg1*; head -c 600 gpurun_out/r2_cfg1_source.csv
