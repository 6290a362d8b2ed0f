#!/bin/bash
# retry while the pod answers busy (exit 3: nothing charged)
for i in 1 2 3 4 5 6 7 8; do
  /usr/local/graft/bin/gpurun --timeout 900 -- 'python -m pytest tests -m gpu -x -q > gpurun_out/s34_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s34_pytest.log; python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s34_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/s34_smoke.log; python bench.py --impl reference > gpurun_out/s34_ref.json 2> gpurun_out/s34_ref.err; python bench.py > gpurun_out/s34_bench.json 2> gpurun_out/s34_bench.err; tail -15 gpurun_out/s34_pytest.log; tail -3 gpurun_out/s34_smoke.log; cut -c1-400 gpurun_out/s34_bench.json'
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 45
done
exit 3
