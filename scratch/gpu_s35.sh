#!/bin/bash
for i in 1 2 3 4 5 6; do
  /usr/local/graft/bin/gpurun --timeout 900 -- 'python -m pytest tests -m gpu -x -q > gpurun_out/s35_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s35_pytest.log; python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s35_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/s35_smoke.log; python bench.py > gpurun_out/s35_bench.json 2> gpurun_out/s35_bench.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s35_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/s35_ncu.log 2>&1; tail -4 gpurun_out/s35_pytest.log; tail -2 gpurun_out/s35_smoke.log; cut -c1-300 gpurun_out/s35_bench.json; wc -l gpurun_out/s35_launches.csv'
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 30
done
exit 3
