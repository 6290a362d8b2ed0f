# Q1/Q6 SF100: bench line with the NVML sampler, ncu launch lists (time + dram bytes), L2 fetch granularity A/B for Q6
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/b2_q1.json 2> gpurun_out/b2_q1.err; echo "bench rc=$?"
for g in 32 128; do
  B200_L2_FETCH=$g timeout 300 python bench.py --device-only q6 --steps 10 --warmup 3 > gpurun_out/b2_q6_l2_$g.json 2>&1; echo "l2 $g rc=$?"
done
timeout 300 python bench.py --device-only q6 --steps 10 --warmup 3 > gpurun_out/b2_q6_l2_default.json 2>&1
cat gpurun_out/b2_q6_l2_*.json
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r2_q1_sf100_launches.csv python bench.py --device-only q1 --steps 1 --warmup 1 > gpurun_out/ncu_q1.log 2>&1; echo "ncu q1 rc=$?"
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r2_q6_sf100_launches.csv python bench.py --device-only q6 --steps 1 --warmup 1 > gpurun_out/ncu_q6.log 2>&1; echo "ncu q6 rc=$?"
tail -2 gpurun_out/ncu_q1.log gpurun_out/ncu_q6.log; tail -c 400 gpurun_out/b2_q1.err
