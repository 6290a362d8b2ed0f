mkdir -p gpurun_out
(nproc; free -g | head -2; nvidia-smi -L | head -3) > gpurun_out/box.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_comm.py -x -q -m gpu > gpurun_out/t_comm.log 2>&1; echo "comm rc=$?" >> gpurun_out/box.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/b_q1.json 2> gpurun_out/b_q1.err; echo "bench rc=$?" >> gpurun_out/box.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/b_ref.json 2> gpurun_out/b_ref.err; echo "ref rc=$?" >> gpurun_out/box.txt
tail -3 gpurun_out/t_comm.log; cat gpurun_out/box.txt; tail -c 600 gpurun_out/b_q1.err
