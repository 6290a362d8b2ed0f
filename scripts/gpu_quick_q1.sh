python -m pytest tests/test_gpu_calc_aggr.py tests/test_gpu_group.py -x -q -m gpu 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu --no-nil-run > gpurun_out/bq.json 2>gpurun_out/bq.err
python -c "
import json
d=json.load(open('gpurun_out/bq.json'))
print('q1', d['ms_per_step'], 'q6', d['config']['q6']['ms_per_step'])
print({k:(round(v['ms_per_launch'],3),round(v['algorithmic_bytes_per_launch']/v['ms_per_launch']/1e6)) for k,v in d['config']['kernels'].items()})
"; tail -c 300 gpurun_out/bq.err
