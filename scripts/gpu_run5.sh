# N GPUs (from $1): strong-scaled bench, repartitioned join
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/b5_q1_n$N.json 2> gpurun_out/b5_q1_n$N.err; echo "bench n$N rc=$?"
tail -c 600 gpurun_out/b5_q1_n$N.err
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/b5_q1_n$N.json').read().strip().splitlines()[-1])
    print('q1', d['ms_per_step'], 'q6', d['config']['q6']['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'q6e2e', d['config']['q6']['e2e']['ms_per_step'], d['config']['nil_run'], d.get('golden_checked'))
    print({k:v['ms_per_step'] for k,v in d['config']['kernels'].items()})
except Exception as e: print('parse failed', e)
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload join --steps 5 --warmup 3 --phases > gpurun_out/b5_join_n$N.json 2> gpurun_out/b5_join_n$N.err; echo "join n$N rc=$?"
tail -c 400 gpurun_out/b5_join_n$N.err; cat gpurun_out/b5_join_n$N.json | cut -c1-1600
