# final N-GPU evidence: strong-scaled contract line, repartitioned join, window (config 5: 1 G rows at N=8)
N=${1:-8}
WROWS=${2:-125000000}
mkdir -p gpurun_out
run() { t=$1; shift; p=$1; shift; timeout $t python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $p bench.py --gpus $N "$@"; }
run 900 29541 --steps 20 --warmup 5 > gpurun_out/f_q1_n$N.json 2> gpurun_out/f_q1_n$N.err; echo "bench n$N rc=$?"
run 600 29542 --workload join --steps 5 --warmup 3 --phases > gpurun_out/f_join_n$N.json 2> gpurun_out/f_join_n$N.err; echo "join n$N rc=$?"
run 900 29543 --workload window --rows $WROWS --steps 3 --warmup 3 > gpurun_out/f_window_n$N.json 2> gpurun_out/f_window_n$N.err; echo "window n$N rc=$?"
python - <<PY
import json
def last(f):
    return json.loads(open(f).read().strip().splitlines()[-1])
try:
    d=last('gpurun_out/f_q1_n$N.json'); c=d['config']
    print('Q1', round(d['ms_per_step'],3), '%.3g'%d['value'], 'q6', round(c['q6']['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],1), 'nil', c['nil_run'] and round(c['nil_run']['q1_ms_per_step'],2), d.get('golden_checked'), d['clocks'])
except Exception as e: print('q1 parse failed', e)
for w in ('join','window'):
    try:
        x=last(f'gpurun_out/f_{w}_n$N.json'); cc=x['config']
        print(w, round(x['ms_per_step'],3), '%.3g'%x['value'], cc['info'].get('phases_ms_last_step'), cc['kernel_ms'])
    except Exception as e: print(w, 'parse failed', e)
PY
tail -c 300 gpurun_out/f_q1_n$N.err gpurun_out/f_join_n$N.err gpurun_out/f_window_n$N.err
