mkdir -p gpurun_out
for ch in default 8 32; do
  if [ $ch != default ]; then export NCCL_MIN_P2P_NCHANNELS=$ch NCCL_MAX_P2P_NCHANNELS=64; fi
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --workload join --steps 5 --warmup 3 --phases 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']
print('$ch', round(d['ms_per_step'],3), c['info']['phases_ms_last_step'], c['kernel_ms'])"
done
