# ncu --set full of the Q1 step's main kernels at the per-GPU size of the 8-GPU run (75 M rows); the report is
# too large to travel back, so the per-kernel detail pages are exported as csv on the box
mkdir -p gpurun_out
REP=/tmp/r2_q1_75m_full
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'aggr_reg_fast|compact_gather|calc_vec|gid8_count|grp_byte_assign' \
  --launch-skip 20 --launch-count 20 -o $REP -f python bench.py --device-only q1 --rows 75004738 --steps 1 --warmup 1 > gpurun_out/ncu_full.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_full.log
ncu -i $REP.ncu-rep --page details --csv > gpurun_out/r2_q1_75m_full_details.csv 2>/dev/null
ncu -i $REP.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__occupancy_limit_registers,smsp__average_warp_latency_issue_stalled_long_scoreboard.pct,dram__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct,launch__grid_size,launch__block_size > gpurun_out/r2_q1_75m_full_raw.csv 2>/dev/null
ls -la gpurun_out/*.csv | tail -3
