# 2 GPUs: NCCL parity test, strong-scaled bench, repartitioned join
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo2.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_comm.py -x -q -m gpu -k nccl > gpurun_out/t_nccl.log 2>&1; echo "nccl test rc=$?"; tail -15 gpurun_out/t_nccl.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/b4_q1_n2.json 2> gpurun_out/b4_q1_n2.err; echo "bench n2 rc=$?"
tail -c 1500 gpurun_out/b4_q1_n2.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/b4_q1_n2.json').read().strip().splitlines()[-1])
    print('q1', d['ms_per_step'], 'q6', d['config']['q6']['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['config']['nil_run'], d.get('golden_checked'))
    print(d['result'])
except Exception as e: print('parse failed', e)
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload join --steps 5 --warmup 3 --phases > gpurun_out/b4_join_n2.json 2> gpurun_out/b4_join_n2.err; echo "join n2 rc=$?"
tail -c 800 gpurun_out/b4_join_n2.err; cat gpurun_out/b4_join_n2.json | cut -c1-1500
