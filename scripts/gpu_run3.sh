mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/t_all.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/b3_q1.json 2> gpurun_out/b3_q1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/b3_q1.json'))
print('q1', d['ms_per_step'], 'q6', d['config']['q6']['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['config']['nil_run'])
print(d['config']['timed_kernel_ms_per_step'])
PY
tail -c 300 gpurun_out/b3_q1.err
