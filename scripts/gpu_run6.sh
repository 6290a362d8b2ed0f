mkdir -p gpurun_out
for w in window config1 join; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/b6_$w.json 2> gpurun_out/b6_$w.err; echo "$w rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/b6_$w.json').read().strip().splitlines()[-1])
c=d['config']
print('$w', 'ms/step', round(d['ms_per_step'],3), 'rows/s', '%.3g'%d['value'], 'frac', round(c['pipeline_frac_of_hbm_peak'],3), 'launches', d['gpu_launches'])
print('  kernel_ms', c['kernel_ms']); print('  counts', c['kernel_launch_counts']); print('  host', c['info'].get('host_ms_per_step'))
PY
done
