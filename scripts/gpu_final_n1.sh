# final N=1 evidence: contract bench line (with cpu_baseline + parity), reference arm, ncu launch lists with DRAM bytes
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/f_q1_n1.json 2> gpurun_out/f_q1_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/f_ref.json 2> gpurun_out/f_ref.err; echo "ref rc=$?"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r2_q1_sf100_launches_v2.csv python bench.py --device-only q1 --steps 1 --warmup 1 > gpurun_out/ncu_q1.log 2>&1; echo "ncu q1 rc=$?"
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r2_q6_sf100_launches_v2.csv python bench.py --device-only q6 --steps 1 --warmup 1 > gpurun_out/ncu_q6.log 2>&1; echo "ncu q6 rc=$?"
for w in window config1 join; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/f_${w}_n1.json 2> gpurun_out/f_${w}_n1.err; echo "$w rc=$?"
done
timeout 900 python bench.py --workload window --rows 500000000 --steps 3 --warmup 3 > gpurun_out/f_window500m_n1.json 2> gpurun_out/f_window500m_n1.err; echo "window500m rc=$?"
tail -c 300 gpurun_out/f_q1_n1.err gpurun_out/f_window500m_n1.err
