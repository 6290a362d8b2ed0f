# round-2 closing N=1 evidence with the final code: full -m gpu suite, smoke, contract bench line (+ reference arm), ncu launch list, other configs
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 > gpurun_out/v3_tests.log; cat gpurun_out/v3_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/v3_smoke.log 2>&1; tail -2 gpurun_out/v3_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/v3_q1_n1.json 2> gpurun_out/v3_q1_n1.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/v3_ref.json 2> gpurun_out/v3_ref.err; echo "ref rc=$?"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r2_q1_sf100_launches_v3.csv python bench.py --device-only q1 --steps 1 --warmup 1 > gpurun_out/ncu_q1.log 2>&1; echo "ncu q1 rc=$?"
for w in window config1 join; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/v3_${w}_n1.json 2> gpurun_out/v3_${w}_n1.err; echo "$w rc=$?"
done
timeout 900 python bench.py --workload window --rows 500000000 --steps 3 --warmup 3 --no-cpu > gpurun_out/v3_window500m_n1.json 2> gpurun_out/v3_window500m_n1.err; echo "window500m rc=$?"
tail -c 300 gpurun_out/v3_q1_n1.err gpurun_out/v3_window500m_n1.err
